"""Host-side logic of the grouped weight-gradient queue (ops/native.py::wgrad), run on CPU with the two native
entry points replaced by recorders: what is launched alone, what is grouped, when the queue is flushed and when the
gradient-ready notifications go out."""
import torch

from distributed_pytorch_cookbook_b200.ops import native


class _Recorder:
    def __init__(self, monkeypatch):
        self.events = []
        monkeypatch.setattr(native, "_WGRAD_GROUP", True)
        monkeypatch.setattr(native, "gemm", self._gemm)
        monkeypatch.setattr(native, "call", self._call)
        monkeypatch.setattr(native, "stream", lambda: 0)
        monkeypatch.setattr(native, "grad_ready_hooks", [lambda p: self.events.append(("ready", p.tag))])
        native._wgrad_queue.clear()
        native._wgrad_cb_armed[0] = False

    def _gemm(self, a, b, out, M, N, K, **kw):
        self.events.append(("gemm", M, N, K))
        return out

    def _call(self, name, n, *rest):
        self.events.append((name, n))
        return 0


def _param(tag, *shape):
    p = torch.nn.Parameter(torch.zeros(*shape))
    p.tag = tag
    return p


class _Layer(torch.autograd.Function):
    """identity whose backward queues one weight gradient per (weight, n_out, k_in) given"""

    @staticmethod
    def forward(ctx, x, specs):
        ctx.specs = specs
        return x.clone()

    @staticmethod
    def backward(ctx, dy):
        tok = 4096
        for w, n_out, k_in in ctx.specs:
            dz = torch.zeros(tok, n_out, dtype=torch.bfloat16)
            xin = torch.zeros(tok, k_in, dtype=torch.bfloat16)
            native.wgrad(dz, xin, torch.zeros(n_out, k_in), n_out, k_in, tok, (w,))
        return dy, None


def test_small_weight_gradients_are_grouped_and_flushed_at_the_end_of_backward(monkeypatch):
    rec = _Recorder(monkeypatch)
    ws = [_param(f"w{i}", 2) for i in range(3)]
    x = torch.ones(2, requires_grad=True)
    y = _Layer.apply(x, [(ws[0], 256, 256), (ws[1], 1024, 256), (ws[2], 256, 1024)])
    y.sum().backward()
    # nothing was launched inside the backward functions; the engine callback grouped all three
    assert rec.events == [("b200_gemm_wgrad_group", 3), ("ready", "w0"), ("ready", "w1"), ("ready", "w2")]
    assert not native._wgrad_queue and not native._wgrad_cb_armed[0]


def test_queue_flushes_at_four_and_large_problems_go_out_at_once(monkeypatch):
    rec = _Recorder(monkeypatch)
    ws = [_param(f"w{i}", 2) for i in range(6)]
    x = torch.ones(2, requires_grad=True)
    specs = [(ws[0], 256, 256), (ws[1], 50257, 256),      # 197 tiles: not worth queueing
             (ws[2], 256, 256), (ws[3], 768, 256), (ws[4], 1024, 256), (ws[5], 256, 256)]
    _Layer.apply(x, specs).sum().backward()
    assert rec.events == [("gemm", 50257, 256, 4096), ("ready", "w1"),
                          ("b200_gemm_wgrad_group", 4), ("ready", "w0"), ("ready", "w2"), ("ready", "w3"), ("ready", "w4"),
                          ("gemm", 256, 256, 4096), ("ready", "w5")]          # the straggler: a plain launch at the end


def test_explicit_flush_and_switch_off(monkeypatch):
    rec = _Recorder(monkeypatch)
    w = [_param("a", 2), _param("b", 2)]

    class Two(torch.autograd.Function):
        @staticmethod
        def forward(ctx, x):
            return x.clone()

        @staticmethod
        def backward(ctx, dy):
            z = torch.zeros(512, 256, dtype=torch.bfloat16)
            native.wgrad(z, z, torch.zeros(256, 256), 256, 256, 512, (w[0],))
            native.wgrad(z, z, torch.zeros(256, 256), 256, 256, 512, (w[1],))
            native.flush_wgrads()                      # what a block's entry LayerNorm does
            rec.events.append(("after_flush",))
            return dy

    x = torch.ones(2, requires_grad=True)
    Two.apply(x).sum().backward()
    assert rec.events == [("b200_gemm_wgrad_group", 2), ("ready", "a"), ("ready", "b"), ("after_flush",)]
    rec.events.clear()
    monkeypatch.setattr(native, "_WGRAD_GROUP", False)
    Two.apply(x).sum().backward()
    assert rec.events == [("gemm", 256, 256, 512), ("ready", "a"), ("gemm", 256, 256, 512), ("ready", "b"), ("after_flush",)]
