TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 4"
F='"ms_per_step": [0-9.]*\|"comm": "[a-z:+]*"'
echo "split+fusedAG";  $TR --master-port 29803 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 6 2>&1 | tail -1 | grep -o "$F" | tr '\n' ' '; echo
echo "split, FUSED_AG=0"; B200GPT_FUSED_AG=0 $TR --master-port 29804 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 6 2>&1 | tail -1 | grep -o "$F" | tr '\n' ' '; echo
echo "rs fused (old), FUSED_AG=0"; B200GPT_FSDP_RS=fused B200GPT_FUSED_AG=0 $TR --master-port 29806 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 6 2>&1 | tail -1 | grep -o "$F" | tr '\n' ' '; echo
echo "nccl"; $TR --master-port 29805 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 6 --comm nccl 2>&1 | tail -1 | grep -o "$F" | tr '\n' ' '; echo
