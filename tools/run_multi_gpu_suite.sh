# Everything that needs N GPUs of one box, results into gpurun_out/ (copy what should be kept to profiles/):
#   bash tools/run_multi_gpu_suite.sh N
# bench.py at N GPUs carries the DDP headline, the self-check and the FSDP 1.4 B / pipeline / pipeline x DDP recipes.
N=$1
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
timeout 400 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 2>&1 | grep "^{" > gpurun_out/bench_${N}gpu_ours.json
cut -c1-200 gpurun_out/bench_${N}gpu_ours.json
timeout 300 $TR --master-port 29541 tools/multi_gpu_check.py 2>&1 | grep "^{" > gpurun_out/multi_gpu_check_${N}gpu.jsonl
grep -c '"ok": true' gpurun_out/multi_gpu_check_${N}gpu.jsonl; grep '"ok": false' gpurun_out/multi_gpu_check_${N}gpu.jsonl | cut -c1-400
timeout 200 $TR --master-port 29542 tools/race_stress.py 300 2>&1 | grep "^{" > gpurun_out/race_stress_${N}gpu.jsonl
cut -c1-200 gpurun_out/race_stress_${N}gpu.jsonl
if [ $N -ge 4 ]; then bash tools/fsdp_ablation.sh $N > gpurun_out/fsdp_ablation_${N}gpu.txt 2>&1; cat gpurun_out/fsdp_ablation_${N}gpu.txt; fi
