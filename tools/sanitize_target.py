"""Small end-to-end target for compute-sanitizer: one eager training step of a tiny GPT (every kernel of the
library launches at least once: embedding, LayerNorm, single-CTA and CTA-pair GEMMs with each epilogue, attention
fwd/bwd, LM head + CE, ReLU backward, AdamW) plus the allocation-free entry point."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from distributed_pytorch_cookbook_b200 import TransformerDecoderLM  # noqa: E402
from distributed_pytorch_cookbook_b200.optim import build_optimizer  # noqa: E402

dev = torch.device("cuda")
torch.manual_seed(0)
B, S, V = 3, 200, 1003          # 597 tokens: M > 128 -> CTA pairs where N >= 256; tails everywhere
model = TransformerDecoderLM(256, 32, 8, 2, V, S).to(dev)
opt = build_optimizer(model, lr=1e-4, device=dev)
ids = torch.randint(0, V, (B, S - 1), device=dev)
pos = torch.arange(S - 1, device=dev).unsqueeze(0).repeat(B, 1)
for _ in range(2):
    opt.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss, _, _ = model.loss(ids, pos, ids)
    loss.backward()
    opt.step()
torch.cuda.synchronize()
print("loss", float(loss.detach()))
