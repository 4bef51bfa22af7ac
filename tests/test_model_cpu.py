"""T1: model parity with the repaired reference on CPU."""
import pytest
import torch

from distributed_pytorch_cookbook_b200 import TransformerDecoderLM, ops
from distributed_pytorch_cookbook_b200.utils import (canonical_state_dict, load_checkpoint,
                                                     prepare_batch, prepare_batch_device,
                                                     save_checkpoint, strip_prefixes)
from oracle import load_reference_gpt, load_reference_utils

CFG = dict(dim=64, head_dim=16, heads=4, num_layers=2, vocab_size=311, max_position_embeddings=32)
ref_gpt = load_reference_gpt()
ref_utils = load_reference_utils()
needs_ref = pytest.mark.skipif(ref_gpt is None, reason="reference source not mounted")


def _pair():
    torch.manual_seed(0)
    ours = TransformerDecoderLM(**CFG)
    ref = ref_gpt.TransformerDecoderLM(**CFG)
    ref.load_state_dict(ours.state_dict())
    return ours, ref


def test_default_config_param_count_and_keys():
    m = TransformerDecoderLM(256, 32, 8, 8, 50257, 256)
    assert sum(p.numel() for p in m.parameters()) == 32_109_568
    sd = m.state_dict()
    assert len(sd) == 109
    assert sd["embeddings.input_embeddings.weight"].shape == (50257, 256)
    assert sd["decoder.layers.7.attn.to_q.weight"].shape == (256, 256)
    assert "decoder.layers.0.attn.to_q.bias" not in sd
    assert sd["decoder.layers.0.fc.up_proj.weight"].shape == (1024, 256)
    assert sd["lm_head.weight"].shape == (50257, 256)


@needs_ref
def test_state_dict_round_trip_both_ways():
    ours, ref = _pair()
    assert list(ours.state_dict().keys()) == list(ref.state_dict().keys())
    for (k1, v1), (k2, v2) in zip(ours.state_dict().items(), ref.state_dict().items()):
        assert v1.shape == v2.shape and v1.dtype == v2.dtype
    ours.load_state_dict(ref.state_dict())


@needs_ref
@pytest.mark.parametrize("with_mask", [False, True])
def test_logits_and_grads_match_reference(with_mask):
    ours, ref = _pair()
    torch.manual_seed(1)
    ids = torch.randint(0, CFG["vocab_size"], (3, 20))
    pos = torch.arange(20).unsqueeze(0).repeat(3, 1)
    mask = None
    if with_mask:
        mask = torch.zeros(3, 20, dtype=torch.bool)
        mask[1, 15:] = True
    a, b = ours(ids, pos, mask), ref(ids, pos, mask)
    valid = torch.ones(3, 20, dtype=torch.bool) if mask is None else ~mask
    assert torch.allclose(a[valid], b[valid], atol=1e-5, rtol=1e-5)
    a[valid].square().sum().backward()
    b[valid].square().sum().backward()
    for (n, p), (_, q) in zip(ours.named_parameters(), ref.named_parameters()):
        assert torch.allclose(p.grad, q.grad, atol=2e-4, rtol=1e-4), n


@needs_ref
def test_loss_matches_reference_cross_entropy():
    ours, ref = _pair()
    ids = torch.randint(0, CFG["vocab_size"], (2, 12))
    pos = torch.arange(12).unsqueeze(0).repeat(2, 1)
    tgt = torch.randint(0, CFG["vocab_size"], (2, 12))
    tgt[0, 3] = -100
    loss, n_valid, n_correct = ours.loss(ids, pos, tgt)
    logits = ref(ids, pos).view(-1, CFG["vocab_size"])
    want = torch.nn.functional.cross_entropy(logits, tgt.view(-1), ignore_index=-100)
    assert torch.allclose(loss, want, atol=1e-5)
    assert int(n_valid) == 23
    m = tgt.view(-1) != -100
    assert int(n_correct) == int((logits.argmax(-1)[m] == tgt.view(-1)[m]).sum())


@needs_ref
def test_prepare_batch_parity():
    torch.manual_seed(0)
    batch = {"input_ids": torch.randint(0, 50, (4, 9)), "attention_mask": torch.ones(4, 9, dtype=torch.int64)}
    batch["attention_mask"][2, 6:] = 0
    a, ta = prepare_batch(batch, 2, "cpu")
    b, tb = ref_utils.prepare_batch(batch, 2, "cpu")
    assert torch.equal(ta, tb)
    for k in b:
        assert torch.equal(a[k], b[k]), k
    c, tc = prepare_batch_device(batch["input_ids"], 2, batch["attention_mask"])
    assert torch.equal(tc, tb)
    for k in b:
        assert torch.equal(c[k], b[k]), k


class _Tok:
    eos_token_id = 7

    def __call__(self, texts, truncation=True, max_length=256, return_tensors="pt"):
        return {"input_ids": torch.tensor([[5, 9, 11]])}

    def decode(self, ids, skip_special_tokens=True):
        return " ".join(str(int(i)) for i in ids)


@needs_ref
def test_generate_matches_reference_tokens():
    from distributed_pytorch_cookbook_b200.utils import generate
    ours, ref = _pair()
    ours.eval(), ref.eval()
    assert generate(ours, "x", _Tok(), torch.device("cpu"), max_new_tokens=8) == \
        ref_utils.generate(ref, "x", _Tok(), torch.device("cpu"), max_new_tokens=8)


def test_generate_with_kv_cache_yields_the_same_tokens_and_logits():
    """use_cache=True (per-layer key/value cache, one token per step) against the reference-style
    full-prefix recompute: identical greedy tokens, and step logits equal to the full forward's."""
    from distributed_pytorch_cookbook_b200.utils import generate
    torch.manual_seed(0)
    m = TransformerDecoderLM(**CFG).eval()
    dev = torch.device("cpu")
    assert generate(m, "x", _Tok(), dev, max_new_tokens=12, use_cache=True) == \
        generate(m, "x", _Tok(), dev, max_new_tokens=12, use_cache=False)
    ids = torch.randint(0, CFG["vocab_size"], (2, 9))
    pos = torch.arange(9).unsqueeze(0).repeat(2, 1)
    with torch.no_grad():
        full = m(ids, pos)
    logits, past = m.forward_cached(ids[:, :6], pos[:, :6])
    assert torch.allclose(logits, full[:, :6], atol=1e-5)
    for t in range(6, 9):
        logits, past = m.forward_cached(ids[:, t:t + 1], pos[:, t:t + 1], past)
        assert torch.allclose(logits[:, 0], full[:, t], atol=1e-5)
    assert past[0][0].shape == (2, CFG["heads"], 9, CFG["head_dim"])


def test_checkpoint_round_trip_and_prefix_stripping(tmp_path):
    torch.manual_seed(0)
    m = TransformerDecoderLM(**CFG)
    path = save_checkpoint(m, tmp_path)
    assert path.name.startswith("checkpoint-") and path.suffix == ".pt"
    m2 = TransformerDecoderLM(**CFG)
    load_checkpoint(m2, path)
    for a, b in zip(m.parameters(), m2.parameters()):
        assert torch.equal(a, b)
    # files written by the reference's compiled / DDP-wrapped models carry prefixes
    pref = {"_orig_mod.module." + k: v for k, v in canonical_state_dict(m).items()}
    torch.save(pref, tmp_path / "ref.pt")
    m3 = TransformerDecoderLM(**CFG)
    load_checkpoint(m3, tmp_path / "ref.pt")
    assert torch.equal(m3.lm_head.weight, m.lm_head.weight)
    assert list(strip_prefixes(pref).keys()) == list(m.state_dict().keys())
    if ref_gpt is not None:   # and the reference can load what we save
        r = ref_gpt.TransformerDecoderLM(**CFG)
        r.load_state_dict(torch.load(path))


def test_attention_reference_is_causal():
    torch.manual_seed(0)
    q, k, v = (torch.randn(1, 2, 6, 4) for _ in range(3))
    o1 = ops.attention_reference(q, k, v, 0.5)
    k2, v2 = k.clone(), v.clone()
    k2[:, :, 4:], v2[:, :, 4:] = 9.0, -9.0
    o2 = ops.attention_reference(q, k2, v2, 0.5)
    assert torch.allclose(o1[:, :, :4], o2[:, :, :4], atol=1e-6)
