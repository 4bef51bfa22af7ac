"""T2: DDP engine over gloo (2 processes) - bucket plan, hook-driven reduction, parity with
single-process training on the concatenated batch (W10-W12)."""
from types import SimpleNamespace

import torch

from dist_utils import run_distributed
from distributed_pytorch_cookbook_b200.parallel.ddp import plan_buckets

CFG = dict(dim=32, head_dim=8, heads=4, num_layers=2, vocab_size=97, max_position_embeddings=16)
ARGS = dict(learning_rate=1e-2, disable_amp=True, bucket_mb=0.01, comm="auto", disable_cuda_graph=True)


def test_bucket_plan_covers_arena_in_reverse_order():
    from distributed_pytorch_cookbook_b200 import TransformerDecoderLM
    from distributed_pytorch_cookbook_b200.runtime.arena import ParamArena
    m = TransformerDecoderLM(**CFG)
    a = ParamArena(list(m.parameters()), "cpu", with_bf16=False, with_moments=False)
    buckets = plan_buckets(a.params, a.offsets, a.total, cap_elems=4096)
    assert buckets[0][1] == a.total and buckets[-1][0] == 0
    for (lo, hi, idxs), nxt in zip(buckets, buckets[1:] + [None]):
        assert lo < hi and idxs == sorted(idxs, reverse=True)
        if nxt is not None:
            assert nxt[1] == lo                     # contiguous, walking towards offset 0
    seen = sorted(i for _, _, idxs in buckets for i in idxs)
    assert seen == list(range(len(a.params)))
    # the default model forms per-layer sized buckets with the reference's 2-bucket tail removed
    big = TransformerDecoderLM(256, 32, 8, 8, 50257, 256)
    ab = ParamArena(list(big.parameters()), "cpu", with_bf16=False, with_moments=False)
    bb = plan_buckets(ab.params, ab.offsets, ab.total, cap_elems=(4 << 20) // 4)
    assert len(bb) >= 10 and bb[0][2] == [len(ab.params) - 1]      # lm_head alone, ready first


def _ddp_worker(rank, world):
    from distributed_pytorch_cookbook_b200 import TransformerDecoderLM
    from distributed_pytorch_cookbook_b200.parallel.ddp import DDPStrategy
    torch.manual_seed(0)
    model = TransformerDecoderLM(**CFG)
    strat = DDPStrategy()
    strat.setup(model, SimpleNamespace(**ARGS), "cpu")
    assert strat.comm_kind == "gloo" and len(strat.reducer.buckets) > 2
    g = torch.Generator().manual_seed(1)
    ids_all = torch.randint(0, CFG["vocab_size"], (3, 4 * world, 12), generator=g)
    pos = torch.arange(12).unsqueeze(0).repeat(4, 1)
    losses = []
    for step in range(3):
        ids = ids_all[step, rank * 4:(rank + 1) * 4]
        losses.append(float(strat.train_step({"input_ids": ids, "position_ids": pos}, ids, amp=False)))
    t = torch.tensor([float(rank), 1.0])
    strat.all_reduce_mean(t)
    return {k: v.numpy().copy() for k, v in strat.state_dict().items()}, losses, t.tolist()


def test_ddp_matches_single_process_on_concatenated_batch():
    world = 2
    outs = run_distributed(_ddp_worker, world)
    sd0, sd1 = outs[0][0], outs[1][0]
    for k in sd0:
        assert (sd0[k] == sd1[k]).all(), k                    # replicas stay in lock-step
    assert outs[0][2] == [0.5, 1.0]
    from distributed_pytorch_cookbook_b200 import TransformerDecoderLM
    torch.manual_seed(0)
    ref = TransformerDecoderLM(**CFG)
    opt = torch.optim.AdamW(ref.parameters(), lr=ARGS["learning_rate"])
    g = torch.Generator().manual_seed(1)
    ids_all = torch.randint(0, CFG["vocab_size"], (3, 4 * world, 12), generator=g)
    pos = torch.arange(12).unsqueeze(0).repeat(4 * world, 1)
    for step in range(3):
        opt.zero_grad()
        loss, _, _ = ref.loss(ids_all[step], pos, ids_all[step])
        loss.backward()
        opt.step()
    for k, v in ref.state_dict().items():
        assert torch.allclose(torch.from_numpy(sd0[k]), v, atol=2e-5, rtol=1e-4), k
