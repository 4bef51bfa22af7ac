"""T1/T2: pipeline partitioner (reference split rule), schedule action lists (deadlock-freedom by
simulation), and the multi-process GPipe / 1F1B / pp x dp executors over gloo vs single-process
training (W17-W19)."""
from types import SimpleNamespace

import pytest
import torch

from dist_utils import run_distributed
from distributed_pytorch_cookbook_b200.parallel.pipe import partition_layers, schedule_actions, stage_layer_range

CFG = dict(dim=32, head_dim=8, heads=4, num_layers=5, vocab_size=101, max_position_embeddings=16)


def test_partition_rule_matches_reference_intent():
    # reference main-pipe.py:57-65: first q+(r==S-1), middle i: q+(i<r), last q
    assert partition_layers(8, 4) == [2, 2, 2, 2]
    assert partition_layers(8, 2) == [4, 4]
    assert partition_layers(8, 3) == [3, 3, 2]          # q=2 r=2=S-1 -> first 3, middle 0 (<2) 3, last 2
    assert partition_layers(5, 4) == [1, 2, 1, 1]       # q=1 r=1: middle 0 gets the extra layer
    assert partition_layers(7, 4) == [2, 2, 2, 1]       # q=1 r=3=S-1
    assert partition_layers(24, 8) == [3] * 8
    assert partition_layers(3, 1) == [3]
    for L in range(1, 30):
        for S in range(1, min(L, 8) + 1):
            per = partition_layers(L, S)
            assert sum(per) == L and len(per) == S
            assert per[-1] == L // S if S > 1 else True   # last stage is the lightest
    assert [stage_layer_range(5, 4, s) for s in range(4)] == [(0, 1), (1, 3), (3, 4), (4, 5)]


@pytest.mark.parametrize("kind", ["gpipe", "1f1b"])
@pytest.mark.parametrize("S,M", [(2, 2), (4, 4), (4, 8), (3, 5), (8, 8), (4, 2)])
def test_schedules_are_complete_and_deadlock_free(kind, S, M):
    acts = [schedule_actions(kind, s, S, M) for s in range(S)]
    for s in range(S):
        assert sorted(a for a in acts[s] if a[0] == "F") == [("F", i) for i in range(M)]
        assert sorted(a for a in acts[s] if a[0] == "B") == [("B", i) for i in range(M)]
        if kind == "1f1b":   # at most S - s micro-batches in flight (the memory bound of 1F1B)
            live = peak = 0
            for a in acts[s]:
                live += 1 if a[0] == "F" else -1
                peak = max(peak, live)
            assert peak <= min(S - s, M)
    # simulate: an F needs the previous stage's F of that micro-batch, a B the next stage's B
    done, pc, progress = set(), [0] * S, True
    while progress:
        progress = False
        for s in range(S):
            if pc[s] == len(acts[s]):
                continue
            kind_, mb = acts[s][pc[s]]
            ready = (s == 0 or ("F", s - 1, mb) in done) if kind_ == "F" else \
                (("F", s, mb) in done and (s == S - 1 or ("B", s + 1, mb) in done))
            if ready:
                done.add((kind_, s, mb)); pc[s] += 1; progress = True
    assert all(pc[s] == len(acts[s]) for s in range(S)), "schedule deadlocks"


def _args(**kw):
    base = dict(learning_rate=1e-2, disable_amp=True, bucket_mb=0.01, comm="auto", disable_cuda_graph=True,
                batch_size=8, sequence_length=13)
    base.update(kw)
    return SimpleNamespace(**base)


def _data(steps, batch):
    g = torch.Generator().manual_seed(7)
    ids = torch.randint(0, CFG["vocab_size"], (steps, batch, 12), generator=g)
    pos = torch.arange(12).unsqueeze(0).repeat(batch, 1)
    return ids, pos


def _pipe_worker(rank, world, schedule, stages, dp, microbatches, checkpoint="never"):
    from distributed_pytorch_cookbook_b200 import TransformerDecoderLM
    from distributed_pytorch_cookbook_b200.parallel.pipe import PipeStrategy
    torch.manual_seed(0)
    model = TransformerDecoderLM(**CFG)
    strat = PipeStrategy(num_stages=stages, num_microbatches=microbatches, schedule=schedule, dp_size=dp,
                         checkpoint=checkpoint)
    strat.setup(model, _args(), "cpu")
    ids, pos = _data(3, 8 * dp)
    losses = []
    r = rank % dp
    for step in range(3):
        x = ids[step, r * 8:(r + 1) * 8]
        losses.append(float(strat.train_step({"input_ids": x, "position_ids": pos[:8]}, x, amp=False)))
    ev = strat.eval_step({"input_ids": ids[0, :8], "position_ids": pos[:8]}, ids[0, :8], amp=False)
    sd = strat.state_dict()
    gen = strat.generation_model()          # collective; rank 0 gets the assembled model
    if rank == 0:
        assert gen is not None
        for k, v in gen.state_dict().items():
            assert torch.equal(v.cpu(), sd[k]), k
        with torch.no_grad():
            assert gen(ids[0, :2], pos[:2]).shape == (2, 12, CFG["vocab_size"])
    else:
        assert gen is None
    return ({k: v.numpy().copy() for k, v in sd.items()} if sd is not None else None, losses,
            [float(e) for e in ev])


def _single_reference(dp):
    from distributed_pytorch_cookbook_b200 import TransformerDecoderLM
    torch.manual_seed(0)
    ref = TransformerDecoderLM(**CFG)
    opt = torch.optim.AdamW(ref.parameters(), lr=1e-2)
    ids, pos = _data(3, 8 * dp)
    losses = []
    for step in range(3):
        opt.zero_grad()
        loss, _, _ = ref.loss(ids[step], pos, ids[step])
        loss.backward()
        opt.step()
        losses.append(float(loss))
    with torch.no_grad():
        ev = ref.loss(ids[0, :8], pos[:8], ids[0, :8])
    return ref, losses, [float(e) for e in ev]


@pytest.mark.parametrize("schedule,stages,dp,mb,ckpt", [("gpipe", 2, 1, 2, "never"), ("1f1b", 4, 1, 4, "never"),
                                                          ("1f1b", 2, 2, 4, "never"), ("1f1b", 3, 1, 4, "except_last"),
                                                          ("gpipe", 2, 1, 4, "always")])
def test_pipeline_matches_single_process(schedule, stages, dp, mb, ckpt):
    """GPipe / 1F1B, more micro-batches than stages, pp x dp, and activation checkpointing (forward re-run inside the
    backward) all train exactly like one process; rank 0 can assemble the full model for generation."""
    outs = run_distributed(_pipe_worker, stages * dp, schedule, stages, dp, mb, ckpt)
    ref, ref_losses, ref_eval = _single_reference(dp)
    sd = outs[0][0]
    assert list(sd.keys()) == list(ref.state_dict().keys())
    for k, v in ref.state_dict().items():
        assert torch.allclose(torch.from_numpy(sd[k]), v, atol=3e-3, rtol=1e-3), k   # Adam amplifies sum-order noise of near-zero grads
    for a, b in zip(outs[0][1], ref_losses):
        assert abs(a - b) < 1e-4
    assert abs(outs[0][2][0] - ref_eval[0]) < 1e-4 and outs[0][2][1] == ref_eval[1]
