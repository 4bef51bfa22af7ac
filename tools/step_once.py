"""Run a few eager (un-graphed) training steps of the flagship config - the command that gets
wrapped in ncu to list every kernel launch of one step with its device time."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from distributed_pytorch_cookbook_b200 import TransformerDecoderLM  # noqa: E402
from distributed_pytorch_cookbook_b200.optim import build_optimizer  # noqa: E402

import os  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
LAYERS = int(os.environ.get("LAYERS", 8))      # LAYERS=1 keeps an ncu --set full capture of one step small
B, S, V = 64, 256, 50257
dev = torch.device("cuda")
torch.manual_seed(0)
model = TransformerDecoderLM(256, 32, 8, LAYERS, V, S).to(dev)
opt = build_optimizer(model, lr=1e-4, device=dev)
ids = torch.randint(0, V, (B, S - 1), device=dev)
pos = torch.arange(S - 1, device=dev).unsqueeze(0).repeat(B, 1)
tgt = torch.randint(0, V, (B, S - 1), device=dev)
for i in range(steps):
    torch.cuda.nvtx.range_push(f"step{i}")
    opt.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss, _, _ = model.loss(ids, pos, tgt)
    loss.backward()
    opt.step()
    torch.cuda.nvtx.range_pop()
torch.cuda.synchronize()
print("loss", float(loss))
