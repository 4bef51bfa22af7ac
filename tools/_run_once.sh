mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -x -q 2>&1 | grep -v Warning | tail -12 > gpurun_out/gpu_tests.txt
cat gpurun_out/gpu_tests.txt
timeout 300 python bench.py --steps 20 --warmup 5 --no-recipes > gpurun_out/bench_grp.json 2> gpurun_out/bench.err
B200GPT_WGRAD_GROUP=0 timeout 300 python bench.py --steps 20 --warmup 5 --no-recipes > gpurun_out/bench_nogrp.json 2>> gpurun_out/bench.err
for f in gpurun_out/bench_grp.json gpurun_out/bench_nogrp.json; do python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['ms_per_step'], d['gpu_launches'])"; done; tail -n 5 gpurun_out/bench.err
