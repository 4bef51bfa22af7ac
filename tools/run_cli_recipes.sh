#!/bin/bash
# End-to-end check of the five recipe scripts on N GPUs (N >= 2; 4 for pipe-ddp): a few synthetic steps each,
# then every checkpoint they wrote is loaded into the REFERENCE's own model class (baseline/_ref).
#   bash tools/run_cli_recipes.sh N
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
COMMON="--synthetic --epochs 1 --max_steps 12 --no_generate"
rm -rf /tmp/ck && mkdir -p /tmp/ck
set -x
python main-single.py $COMMON --checkpoint_dir /tmp/ck/single 2>&1 | grep -a "validation\] Epoch" | tail -1
$TR --master-port 29701 main-ddp.py $COMMON --checkpoint_dir /tmp/ck/ddp 2>&1 | grep -a "validation\] Epoch" | tail -1
$TR --master-port 29702 main-fsdp.py $COMMON --checkpoint_dir /tmp/ck/fsdp 2>&1 | grep -a "validation\] Epoch" | tail -1
$TR --master-port 29703 main-fsdp.py $COMMON --cpu_offload --max_steps 4 --checkpoint_dir /tmp/ck/fsdp_off 2>&1 | grep -a "validation\] Epoch" | tail -1
$TR --master-port 29704 main-pipe.py $COMMON --checkpoint_dir /tmp/ck/pipe 2>&1 | grep -a "validation\] Epoch" | tail -1
if [ "$N" -ge 4 ]; then
  $TR --master-port 29705 main-pipe-ddp.py $COMMON --checkpoint_dir /tmp/ck/pipe_ddp 2>&1 | grep -a "validation\] Epoch" | tail -1
fi
set +x
python - <<'PY'
import glob, sys, torch
sys.path.insert(0, ".")
from baseline.harness import load_reference
ref = load_reference()
ok = True
for f in sorted(glob.glob("/tmp/ck/*/*.pt")):
    m = ref.gpt.TransformerDecoderLM(dim=256, head_dim=32, heads=8, num_layers=8, vocab_size=50257, max_position_embeddings=256)
    try:
        r = m.load_state_dict(torch.load(f, map_location="cpu"))
        print(f, "->", r)
    except Exception as e:
        ok = False
        print(f, "FAILED", e)
print("checkpoints ok" if ok else "CHECKPOINT FAILURE")
PY
