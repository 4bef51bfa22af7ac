TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
N=$1
$TR --nproc-per-node $N --master-port 29601 bench.py --gpus $N --steps 100 --warmup 10 2>&1 | tail -1 > gpurun_out/bench_${N}gpu_ours.json; cut -c1-200 gpurun_out/bench_${N}gpu_ours.json; grep -o '"exposed_comm_ms_per_step": [0-9.]*' gpurun_out/bench_${N}gpu_ours.json
$TR --nproc-per-node $N --master-port 29602 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 8 2>&1 | tail -1 | tee gpurun_out/fsdp_1p3b_${N}gpu.json
$TR --nproc-per-node $N --master-port 29603 tools/bench_recipes.py --recipe fsdp --steps 30 --warmup 8 2>&1 | tail -1 | tee gpurun_out/fsdp_default_${N}gpu.json
if [ $N -ge 4 ]; then $TR --nproc-per-node $N --master-port 29604 tools/bench_recipes.py --recipe pipe-ddp --schedule gpipe --steps 30 --warmup 8 2>&1 | tail -1 | tee gpurun_out/pipe_ddp_${N}gpu.json; fi
if [ $N -le 4 ]; then $TR --nproc-per-node $N --master-port 29605 tools/bench_recipes.py --recipe pipe --schedule gpipe --steps 30 --warmup 8 2>&1 | tail -1 | tee gpurun_out/pipe_${N}gpu.json; fi
$TR --nproc-per-node $N --master-port 29606 tools/multi_gpu_check.py 2>&1 | grep -v "^\[W\|^W1\|^\*\*\*\|OMP_NUM" | tail -12 > gpurun_out/multi_gpu_check_${N}gpu.jsonl; tail -4 gpurun_out/multi_gpu_check_${N}gpu.jsonl | cut -c1-300
