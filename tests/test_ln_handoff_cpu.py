"""Host-side rules of the LayerNorm-backward hand-off (ops/native.py): which producer a LayerNorm may serve under each
scope, and the check that the producer really receives the gradient the LayerNorm handed over."""
import pytest
import torch

from distributed_pytorch_cookbook_b200.ops import native


class _Producer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, kind):
        ctx.fuse_info = (kind, None, 0, None) if kind else None
        return x * 2.0

    @staticmethod
    def backward(ctx, dy):
        ctx.got = native._take_handoff(ctx, dy)
        return dy * 2.0, None


@pytest.fixture(autouse=True)
def _restore_scope():
    yield
    native.set_fuse_scope("all")


def test_scope_rules(monkeypatch):
    monkeypatch.setattr(native, "_LN_FUSE", True)
    x = torch.randn(4, 8, requires_grad=True)
    same_block = _Producer.apply(x, "block")
    next_block = _Producer.apply(x, "stream")
    plain = x * 3.0                                           # not one of our layers
    native.set_fuse_scope("all")
    assert native._producer_of(same_block) == (same_block.grad_fn, True)
    assert native._producer_of(next_block) == (next_block.grad_fn, False)
    assert native._producer_of(plain) == (None, False)
    native.set_fuse_scope("block")                             # FSDP: only inside one transformer block
    assert native._producer_of(same_block) == (same_block.grad_fn, True)
    assert native._producer_of(next_block) == (None, False)
    native.set_fuse_scope("off")
    assert native._producer_of(same_block) == (None, True)     # still known to be mid-block (no wgrad flush there)
    native.set_fuse_scope("all")
    monkeypatch.setattr(native, "_LN_FUSE", False)
    assert native._producer_of(same_block) == (None, True)
    with torch.no_grad():
        assert native._producer_of(same_block) == (None, False)
    assert native._producer_of(same_block.double()) == (None, False)   # fp32 residual stream only


def test_handoff_is_taken_once_and_checked_against_the_incoming_gradient():
    x = torch.randn(4, 8, requires_grad=True)
    y = _Producer.apply(x, "stream")
    node = y.grad_fn
    g = torch.ones(4, 8)
    dz = torch.zeros(4, 8, dtype=torch.bfloat16)
    node.fused = (dz, g.data_ptr())
    y.backward(g)
    assert node.got is dz and not hasattr(node, "fused")
    # a different incoming gradient (autograd summed two consumers into a new tensor): refuse
    y2 = _Producer.apply(x, "stream")
    y2.grad_fn.fused = (dz, g.data_ptr())
    with pytest.raises(RuntimeError, match="second consumer"):
        y2.backward(torch.ones(4, 8))
    # nothing handed over: the layer does its own pass
    y3 = _Producer.apply(x, "stream")
    n3 = y3.grad_fn
    y3.backward(g)
    assert n3.got is None
