mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err
tail -n 1 gpurun_out/bench_2gpu.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d.get('selfcheck')); print(json.dumps(d.get('recipes'), indent=0)[:3000])"
tail -n 5 gpurun_out/bench_2gpu.err
timeout 600 python -m pytest tests -m gpu -x -q -k "multi or ddp or fsdp or pipe or collect" 2>&1 | tail -3
