"""T4/T5: multi-GPU checks (own NVLink collectives vs NCCL, DDP / FSDP / pipeline strategies) -
run tools/multi_gpu_check.py under torchrun on every visible GPU (needs >= 2)."""
import json
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_collectives_and_strategies_on_all_gpus():
    n = torch.cuda.device_count()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29541", str(ROOT / "tools" / "multi_gpu_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT, timeout=900)
    checks = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    bad = [c for c in checks if not c["ok"]]
    assert r.returncode == 0 and checks and not bad, (bad, r.stderr[-3000:])
