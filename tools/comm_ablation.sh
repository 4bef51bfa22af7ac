TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 4"
echo "== DDP default GPT, 4 GPUs: own NVLink all-reduce (symm) vs NCCL"
$TR --master-port 29801 bench.py --gpus 4 --steps 60 --warmup 10 --comm symm 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*\|"comm": "[a-z]*"\|"exposed_comm_ms_per_step": [0-9.]*' | tr '\n' ' '; echo
$TR --master-port 29802 bench.py --gpus 4 --steps 60 --warmup 10 --comm nccl 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*\|"comm": "[a-z]*"\|"exposed_comm_ms_per_step": [0-9.]*' | tr '\n' ' '; echo
echo "== FSDP 1.3B, 4 GPUs: all-gather fused into the GEMM vs side-stream all-gather kernel vs NCCL"
$TR --master-port 29803 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 8 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*\|"comm": "[a-z:+]*"' | tr '\n' ' '; echo
B200GPT_FUSED_AG=0 $TR --master-port 29804 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 8 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*\|"comm": "[a-z:+]*"' | tr '\n' ' '; echo "(B200GPT_FUSED_AG=0)"
$TR --master-port 29805 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 8 --comm nccl 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*\|"comm": "[a-z:+]*"' | tr '\n' ' '; echo
echo "== FSDP default GPT, 4 GPUs"
$TR --master-port 29806 tools/bench_recipes.py --recipe fsdp --steps 30 --warmup 8 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*\|"comm": "[a-z:+]*"' | tr '\n' ' '; echo
B200GPT_FUSED_AG=0 $TR --master-port 29807 tools/bench_recipes.py --recipe fsdp --steps 30 --warmup 8 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*\|"comm": "[a-z:+]*"' | tr '\n' ' '; echo "(B200GPT_FUSED_AG=0)"
$TR --master-port 29808 tools/bench_recipes.py --recipe fsdp --steps 30 --warmup 8 --comm nccl 2>&1 | tail -1 | grep -o '"ms_per_step": [0-9.]*\|"comm": "[a-z:+]*"' | tr '\n' ' '; echo
