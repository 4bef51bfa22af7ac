# FSDP at 1.4 B parameters on N GPUs: own collectives (fused all-gather on CTA pairs / single CTAs / off, one-kernel
# rs_adamw, memory-sharded mode) against the NCCL transport.    bash tools/fsdp_ablation.sh N
N=${1:-4}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
F='"ms_per_step": [0-9.]*\|"comm": "[a-z:+0-9]*"\|"loss_after": [0-9.a-zN]*\|"exposed_comm_ms_per_step": [0-9.]*\|"compute_only_ms_per_step": [0-9.]*'
run() { echo "== $1"; shift; env "$@" timeout 170 $TR --master-port 29803 tools/bench_recipes.py --recipe fsdp --model 1.3b --steps 10 --warmup 6 $EXTRA 2>&1 | grep "^{" | grep -o "$F" | tr '\n' ' '; echo; }
EXTRA="--exposed" run "own collectives, fused all-gather on CTA pairs (default)" X=1
EXTRA="" run "fused all-gather on single CTAs" B200GPT_AG_2SM=0
EXTRA="" run "side-stream all-gather kernel (no fusion)" B200GPT_FUSED_AG=0
EXTRA="" run "one-kernel reduce-scatter + AdamW (round-1 path)" B200GPT_FSDP_RS=fused B200GPT_AG_2SM=0
EXTRA="" run "memory-sharded mode (--reshard_after_forward)" B200GPT_FSDP_RESHARD=1
EXTRA="--comm nccl" run "NCCL transport" X=1
