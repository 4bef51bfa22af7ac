"""T3: every sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

from distributed_pytorch_cookbook_b200.ops import _lib, native

pytestmark = pytest.mark.gpu
BF16 = torch.bfloat16


def _dev_err():
    v = ctypes.c_int(0)
    _lib.call("b200_device_error", ctypes.addressof(v), 1)
    return v.value


@pytest.fixture(autouse=True)
def _need_lib():
    assert _lib.available(), f"native library not loaded: {_lib.load_error()}"
    yield
    torch.cuda.synchronize()
    assert _dev_err() == 0


def test_smoke_launch():
    _lib.call("b200_smoke", _lib.stream())


@pytest.mark.parametrize("n,dim,vocab", [(16320, 256, 50257), (77, 64, 1000)])
def test_embedding_fwd_bwd(n, dim, vocab):
    torch.manual_seed(0)
    ids = torch.randint(0, vocab, (n,), device="cuda")
    pos = torch.arange(n, device="cuda") % 255
    tok = torch.randn(vocab, dim, device="cuda")
    pw = torch.randn(256, dim, device="cuda")
    out = torch.empty(n, dim, device="cuda")
    _lib.call("b200_embedding_fwd", ids.data_ptr(), pos.data_ptr(), tok.data_ptr(), pw.data_ptr(),
              out.data_ptr(), n, dim, vocab, 256, _lib.stream())
    assert torch.equal(out, tok[ids] + pw[pos])
    dout = torch.randn(n, dim, device="cuda")
    dtok, dpos = torch.zeros_like(tok), torch.zeros_like(pw)
    _lib.call("b200_embedding_bwd", ids.data_ptr(), pos.data_ptr(), dout.data_ptr(), dtok.data_ptr(),
              dpos.data_ptr(), n, dim, vocab, 256, _lib.stream())
    rt = torch.zeros_like(tok).index_add_(0, ids, dout)
    rp = torch.zeros_like(pw).index_add_(0, pos, dout)
    assert torch.allclose(dtok, rt, atol=1e-4, rtol=1e-4)
    assert torch.allclose(dpos, rp, atol=1e-3, rtol=1e-4)


@pytest.mark.parametrize("rows,dim", [(16320, 256), (513, 1024), (100, 2048), (33, 192)])
def test_layernorm_fwd_bwd(rows, dim):
    torch.manual_seed(0)
    x = torch.randn(rows, dim, device="cuda") * 2 + 0.5
    g = torch.randn(dim, device="cuda")
    b = torch.randn(dim, device="cuda")
    y = torch.empty(rows, dim, device="cuda", dtype=BF16)
    mean = torch.empty(rows, device="cuda")
    rstd = torch.empty(rows, device="cuda")
    _lib.call("b200_layernorm_fwd", x.data_ptr(), g.data_ptr(), b.data_ptr(), y.data_ptr(),
              mean.data_ptr(), rstd.data_ptr(), rows, dim, 1e-5, _lib.stream())
    xr = x.clone().requires_grad_(True)
    gr, br = g.clone().requires_grad_(True), b.clone().requires_grad_(True)
    ref = F.layer_norm(xr, (dim,), gr, br, 1e-5)
    assert torch.allclose(y.float(), ref, atol=3e-2, rtol=2e-2)
    dy = torch.randn(rows, dim, device="cuda").to(BF16)
    dres = torch.randn(rows, dim, device="cuda")
    ref.backward(dy.float())
    dx = torch.empty(rows, dim, device="cuda")
    dxb = torch.empty(rows, dim, device="cuda", dtype=BF16)
    dg, db = torch.zeros(dim, device="cuda"), torch.zeros(dim, device="cuda")
    _lib.call("b200_layernorm_bwd", dy.data_ptr(), x.data_ptr(), g.data_ptr(), mean.data_ptr(),
              rstd.data_ptr(), dres.data_ptr(), dx.data_ptr(), dxb.data_ptr(), dg.data_ptr(),
              db.data_ptr(), rows, dim, *[t.data_ptr() for t in native.colsum_scratch(x.device)], _lib.stream())
    assert torch.allclose(dx, xr.grad + dres, atol=2e-3, rtol=1e-3)
    assert torch.allclose(dxb.float(), dx, atol=5e-2, rtol=2e-2)
    assert torch.allclose(dg, gr.grad, atol=2e-2 * rows ** 0.5, rtol=1e-3)
    assert torch.allclose(db, br.grad, atol=2e-2 * rows ** 0.5, rtol=1e-3)


def test_adamw_matches_torch():
    torch.manual_seed(0)
    n = 1_000_003
    p = torch.randn(n, device="cuda")
    ref_p = p.clone().requires_grad_(True)
    opt = torch.optim.AdamW([ref_p], lr=1e-3)
    m, v = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    pb = torch.empty(n, device="cuda", dtype=BF16)
    hyper = torch.tensor([1e-3, 0, 0, 0], device="cuda")
    for _ in range(3):
        g = torch.randn(n, device="cuda")
        ref_p.grad = g.clone()
        opt.step()
        _lib.call("b200_adamw_tick", hyper.data_ptr(), _lib.stream())
        _lib.call("b200_adamw", p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), pb.data_ptr(), n,
                  hyper.data_ptr(), 0.9, 0.999, 1e-8, 0.01, _lib.stream())
    assert torch.allclose(p, ref_p.detach(), atol=1e-6, rtol=1e-5)
    assert torch.equal(pb, p.to(BF16))


@pytest.mark.parametrize("fp32", [True, False])
def test_relu_bwd_and_bias_grad(fp32):
    torch.manual_seed(0)
    rows, cols = 1000, 1024
    dy = torch.randn(rows, cols, device="cuda")
    if not fp32:
        dy = dy.to(BF16)
    y = torch.randn(rows, cols, device="cuda").relu().to(BF16)
    dz = torch.empty(rows, cols, device="cuda", dtype=BF16)
    dbias = torch.zeros(cols, device="cuda")
    for _ in range(2):   # twice: the scratch must come back clean
        dbias.zero_()
        _lib.call("b200_relu_bwd", dy.data_ptr(), y.data_ptr(), dz.data_ptr(), dbias.data_ptr(), rows, cols,
                  int(fp32), *[t.data_ptr() for t in native.colsum_scratch(dy.device)], _lib.stream())
    ref = dy.float() * (y.float() > 0)
    assert torch.allclose(dz.float(), ref, atol=2e-2, rtol=1e-2)
    assert torch.allclose(dbias, ref.sum(0), atol=5e-2, rtol=1e-2)


@pytest.mark.parametrize("rows,vocab", [(64, 50257), (33, 1000)])
def test_cross_entropy_fwd_bwd(rows, vocab):
    torch.manual_seed(0)
    ld = (vocab + 127) // 128 * 128
    logits = torch.zeros(rows, ld, device="cuda", dtype=BF16)
    logits[:, :vocab] = (torch.randn(rows, vocab, device="cuda") * 3).to(BF16)
    tgt = torch.randint(0, vocab, (rows,), device="cuda")
    tgt[::7] = -100
    tgt[1] = logits[1, :vocab].float().argmax()
    ref_in = logits[:, :vocab].float().clone().requires_grad_(True)
    ref = F.cross_entropy(ref_in, tgt, ignore_index=-100)
    ref.backward()
    stats = torch.zeros(4, device="cuda")
    _lib.call("b200_count_valid", tgt.data_ptr(), stats.data_ptr(), -100, rows, _lib.stream())
    _lib.call("b200_ce_fwd_bwd", logits.data_ptr(), tgt.data_ptr(), stats.data_ptr(), rows, vocab, ld,
              -100, 1, 1.0, _lib.stream())
    n_valid = int((tgt != -100).sum())
    assert int(stats[0]) == n_valid
    assert abs(float(stats[1]) / n_valid - float(ref)) < 2e-3
    valid = tgt != -100
    want_correct = int((ref_in.argmax(-1)[valid] == tgt[valid]).sum())
    assert int(stats[2]) == want_correct
    assert torch.allclose(logits[:, :vocab].float(), ref_in.grad, atol=2e-3 / n_valid + 1e-4, rtol=3e-2)
    assert float(logits[:, vocab:].abs().max()) == 0.0 if ld > vocab else True


GEMM_SHAPES = [(128, 128, 64), (16320, 768, 256), (1000, 1024, 1024), (300, 50257, 256), (16320, 256, 1024)]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
def test_gemm_forward_epilogues(M, N, K):
    torch.manual_seed(0)
    A = torch.randn(M, K, device="cuda").to(BF16)
    W = (torch.randn(N, K, device="cuda") / K ** 0.5).to(BF16)
    bias = torch.randn(N, device="cuda")
    res = torch.randn(M, N, device="cuda")
    ld = (N + 7) // 8 * 8
    ref = A.float() @ W.float().t()
    out = torch.empty(M, ld, device="cuda", dtype=BF16)
    native.gemm(A, W, out, M, N, K)
    assert torch.allclose(out[:, :N].float(), ref, atol=3e-2, rtol=2e-2)
    if N % 8 == 0:
        out32 = torch.empty(M, N, device="cuda")
        aux = torch.empty(M, N, device="cuda", dtype=BF16)
        native.gemm(A, W, out32, M, N, K, bias=bias, residual=res, act=1, aux_out=aux)
        r2 = (ref + bias).relu()
        assert torch.allclose(aux.float(), r2, atol=3e-2, rtol=2e-2)
        assert torch.allclose(out32, r2 + res, atol=3e-2, rtol=2e-2)


@pytest.mark.parametrize("M,N,K", [(16320, 256, 768), (1024, 256, 50257), (500, 1024, 256)])
def test_gemm_dgrad_mn_major_b(M, N, K):
    """dX[M,N] = dZ[M,K] . W[K,N]  (W stored [K,N]: MN-major B)."""
    torch.manual_seed(0)
    ldk = (K + 7) // 8 * 8                       # TMA needs 16-byte row strides
    dz = torch.zeros(M, ldk, device="cuda", dtype=BF16)
    dz[:, :K] = torch.randn(M, K, device="cuda").to(BF16)
    W = (torch.randn(K, N, device="cuda") / K ** 0.5).to(BF16)
    mask = torch.randn(M, N, device="cuda").to(BF16)
    ref = dz[:, :K].float() @ W.float()
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    native.gemm(dz, W, out, M, N, K, b_mn_major=True)
    assert torch.allclose(out.float(), ref, atol=3e-2, rtol=2e-2)
    native.gemm(dz, W, out, M, N, K, b_mn_major=True, mask_src=mask)
    assert torch.allclose(out.float(), ref * (mask.float() > 0), atol=3e-2, rtol=2e-2)


@pytest.mark.parametrize("Mtok,N,K", [(16320, 1024, 256), (16320, 256, 256), (1024, 50257, 256), (777 * 8, 768, 256)])
def test_gemm_wgrad_mn_major_both_accumulate(Mtok, N, K):
    """dW[N,K] += dZ[Mtok,N]^T . X[Mtok,K] with split-K atomics."""
    torch.manual_seed(0)
    ldz = (N + 7) // 8 * 8
    dzf = torch.zeros(Mtok, ldz, device="cuda")
    dzf[:, :N] = torch.randn(Mtok, N, device="cuda") / Mtok ** 0.5
    dz = dzf.to(BF16)
    X = torch.randn(Mtok, K, device="cuda").to(BF16)
    g0 = torch.randn(N, K, device="cuda")
    g = g0.clone()
    native.gemm(dz, X, g, N, K, Mtok, a_mn_major=True, b_mn_major=True, accumulate=True)
    ref = g0 + dz[:, :N].float().t() @ X.float()
    assert torch.allclose(g, ref, atol=3e-2, rtol=2e-2)


@pytest.mark.parametrize("M,N,K,a_mn,b_mn", [(4096, 1000, 256, False, False), (300, 328, 200, False, False),
                                            (1000, 512, 640, False, True), (1024, 256, 4096, True, True)])
def test_gemm_cta_pair_matches_single_cta(M, N, K, a_mn, b_mn):
    """tcgen05 cta_group::2 (256 x 256 tile on a CTA pair, cfg bit 10) and cta_group::1 (cfg bit 11)
    are both kept alive: each against fp32 PyTorch, and against each other (split-K reduce order
    is the only difference allowed)."""
    torch.manual_seed(0)
    pad = lambda n: (n + 7) // 8 * 8   # noqa: E731
    A = (torch.randn(K, pad(M), device="cuda") if a_mn else torch.randn(M, pad(K), device="cuda")).to(BF16)
    B = (torch.randn(K, pad(N), device="cuda") if b_mn else torch.randn(N, pad(K), device="cuda")).to(BF16)
    Af = (A[:, :M].t() if a_mn else A[:, :K]).float()
    Bf = (B[:, :N].t() if b_mn else B[:, :K]).float()
    base = torch.randn(M, pad(N), device="cuda")
    ref = base[:, :N] + Af @ Bf.t()
    outs = []
    for bit in (0x800, 0x400):
        out = base.clone()
        native.gemm(A, B, out[:, :N], M, N, K, a_mn_major=a_mn, b_mn_major=b_mn, accumulate=True, cfg=bit)
        assert float((out[:, :N] - ref).norm() / ref.norm()) < 1e-3
        outs.append(out)
    assert float((outs[0] - outs[1]).abs().max()) < 1e-3


@pytest.mark.parametrize("M,N,K", [(16320, 1024, 256), (700, 512, 256), (4096, 256, 1024)])
def test_gemm_masked_dgrad_with_fused_bias_gradient(M, N, K):
    """b200_gemm_mask_colsum_bf16: dZ1 = (dZ2 . W2) * (u > 0) and db1 += colsum(dZ1) in one launch."""
    from distributed_pytorch_cookbook_b200.ops.native import call, stream
    torch.manual_seed(0)
    dz2 = torch.randn(M, K, device="cuda").to(BF16)
    W = (torch.randn(K, N, device="cuda") / K ** 0.5).to(BF16)          # stored [K, N]: MN-major B
    u = torch.randn(M, N, device="cuda").to(BF16)
    db0 = torch.randn(N, device="cuda")
    db = db0.clone()
    out = torch.empty(M, N, device="cuda", dtype=BF16)
    for cfg in (0x800, 0x400):                                          # single CTA and CTA pair
        db.copy_(db0)
        call("b200_gemm_mask_colsum_bf16", dz2.data_ptr(), W.data_ptr(), out.data_ptr(), M, N, K, K, N, N, 1,
             u.data_ptr(), db.data_ptr(), cfg, stream())
        ref = (dz2.float() @ W.float()) * (u.float() > 0)
        assert torch.allclose(out.float(), ref, atol=3e-2, rtol=2e-2)
        ref_db = db0 + out.float().sum(0)                               # exactly what the kernel promises
        assert torch.allclose(db, ref_db, atol=2e-2, rtol=2e-3)


def _tiny_model(dev, layers=2, dim=256, heads=8, head_dim=32, vocab=50257, seq=256):
    from distributed_pytorch_cookbook_b200 import TransformerDecoderLM
    torch.manual_seed(0)
    return TransformerDecoderLM(dim=dim, head_dim=head_dim, heads=heads, num_layers=layers,
                                vocab_size=vocab, max_position_embeddings=seq).to(dev)


@pytest.mark.parametrize("B,T,vocab", [(4, 255, 50257), (3, 100, 1000)])
def test_model_loss_and_grads_match_reference_math(B, T, vocab):
    """Whole model through the native kernels vs the same model through PyTorch ops
    (bf16 autocast) and vs fp32: loss and every parameter gradient."""
    from distributed_pytorch_cookbook_b200 import ops
    dev = torch.device("cuda")
    model = _tiny_model(dev, vocab=vocab)
    ids = torch.randint(0, vocab, (B, T), device=dev)
    pos = torch.arange(T, device=dev).unsqueeze(0).repeat(B, 1)
    tgt = torch.randint(0, vocab, (B, T), device=dev)
    tgt[0, :5] = -100
    with torch.autocast("cuda", dtype=BF16):
        loss, n_valid, n_correct = model.loss(ids, pos, tgt)
    loss.backward()
    got = {n: p.grad.clone() for n, p in model.named_parameters()}
    model.zero_grad(set_to_none=True)
    with ops.force_reference():                      # fp32 PyTorch math
        loss32, nv32, _ = model.loss(ids, pos, tgt)
    loss32.backward()
    ref32 = {n: p.grad.clone() for n, p in model.named_parameters()}
    model.zero_grad(set_to_none=True)
    with ops.force_reference(), torch.autocast("cuda", dtype=BF16):   # PyTorch bf16 autocast math
        loss16, _, _ = model.loss(ids, pos, tgt)
    loss16.backward()
    ref16 = {n: p.grad.clone() for n, p in model.named_parameters()}
    assert int(n_valid) == int(nv32) == B * T - 5
    assert abs(float(loss) - float(loss32)) < 3e-2, (float(loss), float(loss32))

    def rel(a, b):
        return ((a - b).norm() / b.norm().clamp(min=1e-12)).item()

    # our bf16 kernels must be as close to the fp32 oracle as PyTorch's own bf16 path is
    worst = {}
    for n in got:
        ours, torch16 = rel(got[n], ref32[n]), rel(ref16[n], ref32[n])
        if ours > max(2.0 * torch16, 0.02):
            worst[n] = (ours, torch16)
    assert not worst, worst


def test_training_reduces_loss_with_fused_adamw_and_graph():
    from distributed_pytorch_cookbook_b200.optim import build_optimizer
    from distributed_pytorch_cookbook_b200.runtime import GraphedStep
    dev = torch.device("cuda")
    model = _tiny_model(dev, vocab=2048)
    opt = build_optimizer(model, lr=1e-3, device=dev)
    ids = torch.randint(0, 2048, (8, 128), device=dev)
    pos = torch.arange(128, device=dev).unsqueeze(0).repeat(8, 1)

    def step(i, p, t):
        opt.zero_grad()
        with torch.autocast("cuda", dtype=BF16):
            loss, _, _ = model.loss(i, p, t)
        loss.backward()
        opt.step()
        return loss.detach()

    g = GraphedStep(step, warmup=3)
    losses = [float(g(ids, pos, ids)) for _ in range(30)]
    assert g.graph is not None
    assert losses[-1] < losses[0] - 1.0, losses


def test_lmhead_ce_chunking_is_invisible(monkeypatch):
    """The fused LM head + CE gives the same loss / counts / gradients whether the rows go through the
    logits buffer in one chunk or in several (B200GPT_LM_CHUNK) or in GEMM+CE sub-chunks (B200GPT_LM_SUB)."""
    torch.manual_seed(0)
    M, Dm, V = 1000, 256, 5003
    x0 = torch.randn(M, Dm, device="cuda").to(BF16)
    w = torch.nn.Parameter(torch.randn(V, Dm, device="cuda") * 0.05)
    tgt = torch.randint(0, V, (M,), device="cuda")
    tgt[:7] = -100
    results = []
    for chunk, sub in ((16384, 0), (384, 0), (512, 128)):
        monkeypatch.setattr(native, "_LM_CHUNK_ROWS", chunk)
        monkeypatch.setattr(native, "_LM_SUB_ROWS", sub)
        x = x0.clone().requires_grad_(True)
        w.grad = None
        loss, n_valid, n_correct = native.lmhead_cross_entropy(x, w, tgt)
        loss.backward()
        results.append((float(loss), int(n_valid), int(n_correct), x.grad.float().clone(), native.grad_of(w).clone()))
    base = results[0]
    for r in results[1:]:
        assert abs(r[0] - base[0]) < 1e-4 and r[1] == base[1] and r[2] == base[2]
        assert torch.allclose(r[3], base[3], atol=2e-3, rtol=2e-2)
        assert torch.allclose(r[4], base[4], atol=1e-4, rtol=1e-3)


def test_lmhead_ce_large_single_chunk_against_fp32():
    """>= 8192 rows in one chunk: dX comes straight out of the un-split dgrad GEMM as bf16."""
    from distributed_pytorch_cookbook_b200 import ops
    torch.manual_seed(0)
    M, Dm, V = 8448, 256, 1003
    x = torch.randn(M, Dm, device="cuda").to(BF16).requires_grad_(True)
    w = torch.nn.Parameter(torch.randn(V, Dm, device="cuda") * 0.05)
    tgt = torch.randint(0, V, (M,), device="cuda")
    loss, n_valid, _ = native.lmhead_cross_entropy(x, w, tgt)
    loss.backward()
    x32 = x.detach().float().requires_grad_(True)
    w32 = w.detach().to(BF16).float().requires_grad_(True)
    ref, nv, _ = ops.lmhead_cross_entropy_reference(x32, w32, tgt)
    ref.backward()
    assert abs(float(loss) - float(ref)) < 2e-3 and int(n_valid) == int(nv)
    assert float((x.grad.float() - x32.grad).norm() / x32.grad.norm()) < 2e-2
    assert float((native.grad_of(w) - w32.grad).norm() / w32.grad.norm()) < 2e-2


@pytest.mark.parametrize("act", ["relu", "gelu", "gelu_tanh"])
@pytest.mark.parametrize("with_res", [True, False])
def test_fused_mlp_activations_against_fp32(act, with_res):
    """The MLP node (two GEMM epilogues, act' folded into the dgrad epilogue) for ReLU / GELU / tanh-GELU against
    fp32 PyTorch on the same bf16-rounded operands: output, dX, dW1, db1, dW2, db2, residual gradient."""
    torch.manual_seed(4)
    M, D, Fd = 1000, 256, 1024
    fn = {"relu": F.relu, "gelu": F.gelu, "gelu_tanh": lambda t: F.gelu(t, approximate="tanh")}[act]
    x = torch.randn(M, D, device="cuda").to(BF16).requires_grad_(True)
    res = torch.randn(M, D, device="cuda").requires_grad_(True) if with_res else None
    w1 = torch.nn.Parameter(torch.randn(Fd, D, device="cuda") * 0.06)
    b1 = torch.nn.Parameter(torch.randn(Fd, device="cuda") * 0.1)
    w2 = torch.nn.Parameter(torch.randn(D, Fd, device="cuda") * 0.03)
    b2 = torch.nn.Parameter(torch.randn(D, device="cuda") * 0.1)
    out = native.mlp(x, w1, b1, w2, b2, residual=res, activation=act)
    dy = torch.randn_like(out, dtype=torch.float32)
    out.float().backward(dy)
    got = [out.float(), x.grad.float(), native.grad_of(w1).clone(), native.grad_of(b1).clone(),
           native.grad_of(w2).clone(), native.grad_of(b2).clone()] + ([res.grad] if with_res else [])
    x32 = x.detach().float().requires_grad_(True)
    p32 = [t.detach().to(BF16).float().requires_grad_(True) if t.ndim == 2 else t.detach().clone().requires_grad_(True)
           for t in (w1, b1, w2, b2)]
    r32 = res.detach().clone().requires_grad_(True) if with_res else None
    h = fn(F.linear(x32, p32[0], p32[1]))
    ref = fn(F.linear(h, p32[2], p32[3]))
    if with_res:
        ref = ref + r32
    ref.backward(dy)
    want = [ref, x32.grad, p32[0].grad, p32[1].grad, p32[2].grad, p32[3].grad] + ([r32.grad] if with_res else [])
    names = ["out", "dx", "dw1", "db1", "dw2", "db2", "dres"]
    # ReLU's derivative is a step: a pre-activation whose bf16 rounding changes sign flips a whole mask bit, which the
    # smooth GELUs do not suffer from
    tol = 6e-2 if act == "relu" else 2.5e-2
    for n, g, w in zip(names, got, want):
        assert _rel(g, w) < tol, (act, n, _rel(g, w))


def test_gelu_model_takes_the_native_path():
    """A GELU FeedForward under bf16 autocast runs on the fused kernels (no silent drop to eager PyTorch)."""
    from distributed_pytorch_cookbook_b200.models.gpt import FeedForward
    torch.manual_seed(0)
    ff = FeedForward(256, activation=F.gelu).cuda()
    x = torch.randn(4, 50, 256, device="cuda")
    before = _lib.launch_count()
    with torch.autocast("cuda", dtype=BF16):
        y = ff(x)
    assert _lib.launch_count() - before >= 2
    ref = ff(x)
    assert _rel(y, ref) < 2e-2


def _lm_reference(x, w, tgt):
    from distributed_pytorch_cookbook_b200 import ops
    x32 = x.detach().float().requires_grad_(True)
    w32 = w.detach().to(BF16).float().requires_grad_(True)
    ref, nv, nc = ops.lmhead_cross_entropy_reference(x32, w32, tgt)
    ref.backward()
    return float(ref), int(nv), int(nc), x32.grad, w32.grad


def _rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp(min=1e-20))


@pytest.mark.parametrize("M,Dm,V,xs", [(16320, 256, 50257, 1.0), (1000, 256, 5003, 1.0), (300, 512, 1000, 6.0),
                                         (129, 64, 257, 1.0)])
def test_lmhead_fused_softmax_epilogue_against_fp32(M, Dm, V, xs):
    """Softmax-epilogue LM head (no logits in memory): loss, counts, dX and dW against fp32 PyTorch on the
    same bf16 operands, with ignored rows, a vocabulary tail and logits of very different magnitude."""
    torch.manual_seed(1)
    x = (torch.randn(M, Dm, device="cuda") * xs).to(BF16).requires_grad_(True)
    w = torch.nn.Parameter(torch.randn(V, Dm, device="cuda") * 0.05)
    tgt = torch.randint(0, V, (M,), device="cuda")
    tgt[3:9] = -100
    tgt[-1] = V - 1
    for it in range(2):      # second call runs with the gap observed by the first
        x.grad = None
        w.grad = None
        loss, n_valid, n_correct = native.lmhead_cross_entropy(x, w, tgt)
        loss.backward()
        ref, nv, nc, dx32, dw32 = _lm_reference(x, w, tgt)
        assert abs(float(loss) - ref) < 2e-3 * max(1.0, abs(ref)), (it, float(loss), ref)
        assert int(n_valid) == nv == M - 6
        assert abs(int(n_correct) - nc) <= max(1, nc // 50), (int(n_correct), nc)
        assert _rel(x.grad, dx32) < 2e-2, (it, _rel(x.grad, dx32))
        assert _rel(native.grad_of(w), dw32) < 2e-2, (it, _rel(native.grad_of(w), dw32))
        assert float(x.grad[3:9].abs().max()) == 0.0          # ignored rows get no gradient
    st = native.lm_state(w).tolist()
    assert st[3] == 0.0 and st[1] >= 0.0, st                  # nothing had to be recomputed
    assert _dev_err() == 0


def test_lmhead_rows_outside_the_shift_window_are_recomputed():
    """A wrong gap estimate (here: absurdly large, so every exp overflows) must only cost time: finish
    recomputes the rows exactly, and the next call runs with the gap that was observed."""
    torch.manual_seed(2)
    M, Dm, V = 200, 256, 3001
    x = torch.randn(M, Dm, device="cuda").to(BF16).requires_grad_(True)
    w = torch.nn.Parameter(torch.randn(V, Dm, device="cuda") * 0.05)
    tgt = torch.randint(0, V, (M,), device="cuda")
    tgt[:4] = -100
    ref, nv, nc, dx32, dw32 = _lm_reference(x, w, tgt)
    st = native.lm_state(w)
    for poison, expect_bad in ((900.0, M), (None, 0), (-1.0, 0)):
        if poison is not None and poison > 0:
            st[2] = poison
        x.grad = None
        w.grad = None
        loss, n_valid, n_correct = native.lmhead_cross_entropy(x, w, tgt)
        loss.backward()
        torch.cuda.synchronize()
        assert st[3].item() == expect_bad, (poison, st.tolist())
        assert abs(float(loss) - ref) < 2e-3 * max(1.0, abs(ref)), (poison, float(loss), ref)
        assert abs(int(n_correct) - nc) <= 1 and int(n_valid) == nv
        assert _rel(x.grad, dx32) < 2e-2 and _rel(native.grad_of(w), dw32) < 2e-2
    # evaluation (no gradient, nothing stored) agrees too, also from a poisoned estimate
    st[2] = 900.0
    with torch.no_grad():
        l2, nv2, nc2 = native.lmhead_cross_entropy(x.detach(), w, tgt)
    assert abs(float(l2) - ref) < 2e-3 * max(1.0, abs(ref)) and int(nv2) == nv and abs(int(nc2) - nc) <= 1
    assert _dev_err() == 0


def test_lmhead_fused_and_unfused_paths_agree(monkeypatch):
    torch.manual_seed(3)
    M, Dm, V = 2048, 256, 50257
    x0 = torch.randn(M, Dm, device="cuda").to(BF16)
    w = torch.nn.Parameter(torch.randn(V, Dm, device="cuda") * 0.03)
    tgt = torch.randint(0, V, (M,), device="cuda")
    out = []
    for fused in (True, False):
        monkeypatch.setattr(native, "_LM_FUSED", fused)
        x = x0.clone().requires_grad_(True)
        w.grad = None
        loss, nv, nc = native.lmhead_cross_entropy(x, w, tgt, grad_scale=0.5)
        (loss * 0.5).backward()
        out.append((float(loss), int(nv), int(nc), x.grad.clone(), native.grad_of(w).clone()))
    a, b = out
    assert abs(a[0] - b[0]) < 2e-3 and a[1] == b[1] and abs(a[2] - b[2]) <= 2
    assert _rel(a[3], b[3]) < 2e-2 and _rel(a[4], b[4]) < 2e-2


def test_lmhead_wrong_loss_gradient_is_detected(monkeypatch):
    """The op bakes grad_scale into the gradients it produces in forward; a caller that scales the loss by
    something else must hear about it (first un-captured calls, or always with B200GPT_DEBUG=1)."""
    monkeypatch.setenv("B200GPT_DEBUG", "1")
    x = torch.randn(64, 64, device="cuda").to(BF16).requires_grad_(True)
    w = torch.nn.Parameter(torch.randn(300, 64, device="cuda") * 0.05)
    tgt = torch.randint(0, 300, (64,), device="cuda")
    loss, _, _ = native.lmhead_cross_entropy(x, w, tgt)
    with pytest.raises(RuntimeError, match="grad_scale"):
        (loss * 0.25).backward()


def test_model_forward_returns_logits_at_odd_vocab():
    """Reference-style ``model(input_ids, position_ids)`` (gpt.py:221-231) under bf16 autocast at vocab 50257:
    the logits rows are not a multiple of 16 bytes, the op pads them and returns the [.., vocab] view; its
    backward and greedy generation without the cache go through the same path."""
    from distributed_pytorch_cookbook_b200 import ops
    dev = torch.device("cuda")
    V = 50257
    model = _tiny_model(dev, layers=1, vocab=V)
    ids = torch.randint(0, V, (2, 37), device=dev)
    pos = torch.arange(37, device=dev).unsqueeze(0).repeat(2, 1)
    with torch.autocast("cuda", dtype=BF16):
        logits = model(ids, pos)
    assert logits.shape == (2, 37, V) and logits.dtype == BF16
    with ops.force_reference():
        ref = model(ids, pos)
    assert float((logits.float() - ref).abs().max()) < 0.08 * max(1.0, float(ref.abs().max()))
    tgt = torch.randint(0, V, (2, 37), device=dev)
    with torch.autocast("cuda", dtype=BF16):
        loss = F.cross_entropy(model(ids, pos).float().view(-1, V), tgt.view(-1))
    loss.backward()
    g = model.lm_head.weight.grad.clone()
    model.zero_grad(set_to_none=True)
    with ops.force_reference():
        F.cross_entropy(model(ids, pos).view(-1, V), tgt.view(-1)).backward()
    assert _rel(g, model.lm_head.weight.grad) < 3e-2
    assert _dev_err() == 0


@pytest.mark.parametrize("B,T,H,D,masked", [(2, 255, 8, 32, False), (2, 100, 2, 64, True), (1, 300, 2, 128, False),
                                            (64, 255, 8, 32, False), (1, 1023, 2, 128, True), (2, 1024, 3, 128, False),
                                            (1, 641, 2, 64, True), (1, 1, 2, 32, False), (2, 127, 2, 64, False),
                                            (2, 129, 2, 128, False), (1, 2000, 1, 128, False)])
def test_flash_attention_fwd_bwd(B, T, H, D, masked):
    from distributed_pytorch_cookbook_b200 import ops
    torch.manual_seed(0)
    qkv = (torch.randn(B, T, 3 * H * D, device="cuda") * 0.7).to(BF16).requires_grad_(True)
    mask = None
    if masked:
        mask = torch.zeros(B, T, dtype=torch.bool, device="cuda")
        mask[0, T - 17:] = True
    scale = D ** -0.5
    out = native.attention(qkv, H, D, scale, mask)
    dout = torch.randn_like(out)
    out.backward(dout)
    got_dqkv = qkv.grad.clone()
    q32 = qkv.detach().float().requires_grad_(True)
    q, k, v = (t.reshape(B, T, H, D).transpose(1, 2) for t in q32.chunk(3, dim=-1))
    ref = ops.attention_reference(q, k, v, scale, mask).transpose(1, 2).reshape(B, T, H * D)
    ref.backward(dout.float())
    valid = torch.ones(B, T, dtype=torch.bool, device="cuda") if mask is None else ~mask
    assert torch.allclose(out.float()[valid], ref[valid], atol=3e-2, rtol=3e-2)
    gd = (got_dqkv.float() - q32.grad)[valid]
    assert gd.abs().max().item() <= 0.05 * q32.grad.abs().max().item() + 2e-2, gd.abs().max().item()


@pytest.mark.parametrize("B,T,H,D", [(2, 1023, 4, 128), (1, 513, 2, 64), (3, 255, 2, 32)])
def test_attention_forward_kernels_agree(B, T, H, D):
    """The per-tile forward and the warp-specialised, TMA-fed forward (the default for long sequences) compute the
    same O and log-sum-exp; both are checked against fp32."""
    import ctypes
    from distributed_pytorch_cookbook_b200 import ops
    torch.manual_seed(5)
    qkv = (torch.randn(B, T, 3 * H * D, device="cuda") * 0.8).to(BF16)
    scale = D ** -0.5
    q, k, v = (t.float().reshape(B, T, H, D).transpose(1, 2) for t in qkv.chunk(3, dim=-1))
    ref = ops.attention_reference(q, k, v, scale, None).transpose(1, 2).reshape(B, T, H * D)
    outs = []
    for ws in (0, 1):
        out = torch.empty((B, T, H * D), dtype=BF16, device="cuda")
        lse = torch.empty((B, H, T), dtype=torch.float32, device="cuda")
        _lib.call("b200_attn_fwd_select", qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, T, H, D, float(scale), None,
                  ws, _lib.stream())
        torch.cuda.synchronize()
        assert torch.allclose(out.float(), ref, atol=3e-2, rtol=3e-2), ws
        outs.append((out, lse))
    assert torch.allclose(outs[0][0].float(), outs[1][0].float(), atol=2e-2, rtol=2e-2)
    assert torch.allclose(outs[0][1], outs[1][1], atol=2e-3, rtol=1e-3)
    assert _dev_err() == 0


@pytest.mark.parametrize("B,T,H,D,masked", [(2, 1023, 3, 128, False), (1, 641, 2, 64, True), (1, 2000, 1, 128, False)])
def test_attention_backward_kernels_agree(B, T, H, D, masked):
    """Three backward implementations - per-tile kernel, warp-specialised kernel reducing dQ with fp32 vector reductions,
    warp-specialised dK / dV kernel + query-owned dQ kernel (no atomics) - against fp32 autograd and each other."""
    from distributed_pytorch_cookbook_b200 import ops
    torch.manual_seed(6)
    qkv = (torch.randn(B, T, 3 * H * D, device="cuda") * 0.7).to(BF16)
    scale = D ** -0.5
    mask = None
    km = None
    if masked:
        mask = torch.zeros(B, T, dtype=torch.bool, device="cuda")
        mask[0, T - 29:] = True
        km = mask.to(torch.uint8).contiguous()
    out = torch.empty((B, T, H * D), dtype=BF16, device="cuda")
    lse = torch.empty((B, H, T), dtype=torch.float32, device="cuda")
    _lib.call("b200_attn_fwd", qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, T, H, D, float(scale),
              _lib.ptr(km), _lib.stream())
    dout = torch.randn_like(out)
    q32 = qkv.float().requires_grad_(True)
    q, k, v = (t.reshape(B, T, H, D).transpose(1, 2) for t in q32.chunk(3, dim=-1))
    ops.attention_reference(q, k, v, scale, mask).transpose(1, 2).reshape(B, T, H * D).backward(dout.float())
    valid = torch.ones(B, T, dtype=torch.bool, device="cuda") if mask is None else ~mask
    got = []
    for mode in (0, 1, 2):
        dqkv = torch.full_like(qkv, float("nan"))
        delta = torch.empty((B, H, T), dtype=torch.float32, device="cuda")
        dq_acc = torch.empty((B, T, H * D), dtype=torch.float32, device="cuda")
        _lib.call("b200_attn_bwd_select", qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(),
                  delta.data_ptr(), dq_acc.data_ptr(), B, T, H, D, float(scale), _lib.ptr(km), mode, _lib.stream())
        torch.cuda.synchronize()
        gd = (dqkv.float() - q32.grad)[valid]
        assert gd.abs().max().item() <= 0.05 * q32.grad.abs().max().item() + 2e-2, (mode, gd.abs().max().item())
        got.append(dqkv.float()[valid])
    assert torch.allclose(got[0], got[1], atol=3e-2, rtol=3e-2) and torch.allclose(got[1], got[2], atol=3e-2, rtol=3e-2)
    assert _dev_err() == 0


@pytest.mark.parametrize("dim,heads,act", [(256, 8, "relu"), (512, 8, "gelu"), (192, 6, "relu")])
def test_layernorm_backward_handoff_matches_separate_passes(monkeypatch, dim, heads, act):
    """The LayerNorm backward that also writes the producer layer's dz / bias gradient (b200_layernorm_bwd_fused)
    gives the same gradients as the separate relu_bwd / act_bwd passes, and launches fewer kernels."""
    from distributed_pytorch_cookbook_b200 import TransformerDecoderLM
    from distributed_pytorch_cookbook_b200.ops import native
    dev = torch.device("cuda")
    vocab, B, T = 1000, 3, 127
    torch.manual_seed(0)
    model = TransformerDecoderLM(dim=dim, head_dim=32, heads=heads, num_layers=2, vocab_size=vocab,
                                 max_position_embeddings=T).to(dev)
    if act == "gelu":
        for layer in model.decoder.layers:
            layer.fc.activation = F.gelu
    ids = torch.randint(0, vocab, (B, T), device=dev)
    pos = torch.arange(T, device=dev).unsqueeze(0).repeat(B, 1)
    tgt = torch.randint(0, vocab, (B, T), device=dev)

    def run(fuse):
        monkeypatch.setattr(native, "_LN_FUSE", fuse)
        model.zero_grad(set_to_none=True)
        n0 = _lib.launch_count()
        with torch.autocast("cuda", dtype=BF16):
            loss, _, _ = model.loss(ids, pos, tgt)
        loss.backward()
        torch.cuda.synchronize()
        return float(loss), {n: p.grad.clone() for n, p in model.named_parameters()}, _lib.launch_count() - n0

    run(True)                                # one-time launches (LM-head state, first-call checks) stay out of the count
    l1, g1, k1 = run(True)
    l0, g0, k0 = run(False)
    assert abs(l1 - l0) < 1e-5               # same forward (its atomics make the last digit vary)
    assert k1 == k0 - 4, (k1, k0)            # 2 layers x (out-projection cast + MLP activation backward)
    for n in g0:
        assert _rel(g1[n], g0[n]) < 2e-3, (n, _rel(g1[n], g0[n]))


def test_layernorm_backward_handoff_detects_a_second_consumer():
    """If the residual stream feeds something besides the LayerNorm, the producer's incoming gradient is a sum the
    LayerNorm never saw: the hand-off must fail loudly instead of using the stale dz."""
    from distributed_pytorch_cookbook_b200 import ops
    dev = torch.device("cuda")
    torch.manual_seed(0)
    D = 256
    lin = torch.nn.Linear(D, D).to(dev)
    ln = torch.nn.LayerNorm(D).to(dev)
    x = torch.randn(64, D, device=dev, requires_grad=True)
    res = torch.randn(64, D, device=dev)
    with torch.autocast("cuda", dtype=BF16):
        y = ops.linear(x.to(BF16), lin.weight, lin.bias, residual=res, out_fp32=True)
        _, h = ops.layernorm_residual(y, ln.weight, ln.bias, ln.eps)
        loss = h.float().sum() + y.sum()                 # second consumer of y
    with pytest.raises(RuntimeError, match="second consumer"):
        loss.backward()


@pytest.mark.parametrize("tok,shapes", [(16320, [(768, 256), (256, 256), (1024, 256), (256, 1024)]),
                                        (1000, [(768, 256), (264, 520)]), (4096, [(296, 256), (512, 512), (256, 768)])])
def test_grouped_weight_gradient_gemm(tok, shapes):
    """b200_gemm_wgrad_group: several dW += dZ^T X problems in one launch vs fp32 matmuls (accumulating into a
    non-zero gradient)."""
    import ctypes
    from distributed_pytorch_cookbook_b200.ops.native import call, stream
    dev = torch.device("cuda")
    torch.manual_seed(0)
    dz = [torch.randn(tok, n, device=dev).to(BF16) for n, _ in shapes]
    xs = [torch.randn(tok, k, device=dev).to(BF16) for _, k in shapes]
    init = [torch.randn(n, k, device=dev) for n, k in shapes]
    out = [t.clone() for t in init]
    n = len(shapes)
    vp, ci = ctypes.c_void_p, ctypes.c_int
    arr = lambda t, vals: ctypes.cast((t * n)(*vals), vp)   # noqa: E731
    call("b200_gemm_wgrad_group", n, arr(vp, [t.data_ptr() for t in dz]), arr(vp, [t.data_ptr() for t in xs]),
         arr(vp, [t.data_ptr() for t in out]), arr(ci, [s[0] for s in shapes]), arr(ci, [s[1] for s in shapes]), tok,
         arr(ci, [t.stride(0) for t in dz]), arr(ci, [t.stride(0) for t in xs]), arr(ci, [t.stride(0) for t in out]),
         stream())
    torch.cuda.synchronize()
    for i in range(n):
        ref = init[i] + dz[i].float().t() @ xs[i].float()
        assert _rel(out[i], ref) < 2e-3, (i, shapes[i], _rel(out[i], ref))


def test_grouped_weight_gradients_in_the_model_match_separate_launches(monkeypatch):
    """Deferring and grouping the dW GEMMs of a block changes the launch count, not the gradients."""
    from distributed_pytorch_cookbook_b200.ops import native
    dev = torch.device("cuda")
    model = _tiny_model(dev, vocab=1000, seq=128)
    B, T = 8, 127
    ids = torch.randint(0, 1000, (B, T), device=dev)
    pos = torch.arange(T, device=dev).unsqueeze(0).repeat(B, 1)
    tgt = torch.randint(0, 1000, (B, T), device=dev)

    def run(group):
        monkeypatch.setattr(native, "_WGRAD_GROUP", group)
        model.zero_grad(set_to_none=True)
        n0 = _lib.launch_count()
        with torch.autocast("cuda", dtype=BF16):
            loss, _, _ = model.loss(ids, pos, tgt)
        loss.backward()
        torch.cuda.synchronize()
        assert not native._wgrad_queue
        return {n: p.grad.clone() for n, p in model.named_parameters()}, _lib.launch_count() - n0

    run(True)
    g1, k1 = run(True)
    g0, k0 = run(False)
    assert k1 <= k0 - 2 * 3, (k1, k0)            # 2 layers: at least four dW launches become one
    for n in g0:
        assert _rel(g1[n], g0[n]) < 2e-3, (n, _rel(g1[n], g0[n]))
