"""Property tests (hypothesis) for the host-side planners: DDP bucket plan, pipeline partition rule and
schedules.  Pure Python, no process groups."""
from hypothesis import given, settings, strategies as st

from distributed_pytorch_cookbook_b200.parallel.ddp import plan_buckets
from distributed_pytorch_cookbook_b200.parallel.pipe import partition_layers, schedule_actions, stage_layer_range


@settings(max_examples=200, deadline=None)
@given(sizes=st.lists(st.integers(1, 5000), min_size=1, max_size=40), cap=st.integers(1, 20000))
def test_bucket_plan_partitions_any_arena(sizes, cap):
    """Buckets tile [0, total) exactly, walk from the end of the arena to offset 0, hold every parameter
    once, and only a single over-sized parameter may exceed the cap."""
    sizes = [4 * s for s in sizes]                      # arena slots are 16-byte granular
    offsets, total = [], 0
    for s in sizes:
        offsets.append(total)
        total += s
    buckets = plan_buckets([None] * len(sizes), offsets, total, cap_elems=cap)
    assert buckets[0][1] == total and buckets[-1][0] == 0
    for (lo, hi, idxs), nxt in zip(buckets, buckets[1:] + [None]):
        assert lo < hi and idxs == sorted(idxs, reverse=True)
        assert lo == offsets[idxs[-1]] and hi == offsets[idxs[0]] + sizes[idxs[0]]
        assert hi - lo <= cap or len(idxs) == 1
        if nxt is not None:
            assert nxt[1] == lo
    assert sorted(i for _, _, idxs in buckets for i in idxs) == list(range(len(sizes)))


@settings(max_examples=200, deadline=None)
@given(L=st.integers(1, 64), S=st.integers(1, 16))
def test_partition_rule_properties(L, S):
    if S > L:
        return
    per = partition_layers(L, S)
    assert len(per) == S and sum(per) == L and min(per) >= 1 or L < S
    assert max(per) - min(per) <= 1
    if S > 1:
        assert per[-1] == L // S                        # the stage that also holds the LM head is lightest
    ranges = [stage_layer_range(L, S, s) for s in range(S)]
    assert ranges[0][0] == 0 and ranges[-1][1] == L
    assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))


@settings(max_examples=150, deadline=None)
@given(kind=st.sampled_from(["gpipe", "1f1b"]), S=st.integers(1, 8), M=st.integers(1, 12))
def test_schedules_run_to_completion_with_bounded_activations(kind, S, M):
    """Simulate the per-stage action lists with blocking send/recv semantics: no deadlock, every
    micro-batch forward before its backward on every stage, and 1F1B never holds more than
    min(S - stage, M) activations."""
    acts = [schedule_actions(kind, s, S, M) for s in range(S)]
    pc = [0] * S
    done_f = [set() for _ in range(S)]
    done_b = [set() for _ in range(S)]
    live = [0] * S
    peak = [0] * S
    progressed = True
    while progressed:
        progressed = False
        for s in range(S):
            if pc[s] == len(acts[s]):
                continue
            k, mb = acts[s][pc[s]]
            ready = (k == "F" and (s == 0 or mb in done_f[s - 1])) or \
                    (k == "B" and mb in done_f[s] and (s == S - 1 or mb in done_b[s + 1]))
            if ready:
                (done_f if k == "F" else done_b)[s].add(mb)
                live[s] += 1 if k == "F" else -1
                peak[s] = max(peak[s], live[s])
                pc[s] += 1
                progressed = True
    for s in range(S):
        assert pc[s] == len(acts[s]), f"stage {s} stuck at {acts[s][pc[s]]}"
        assert done_f[s] == done_b[s] == set(range(M)) and live[s] == 0
        if kind == "1f1b":
            assert peak[s] <= min(S - s, M)
