"""Codec numerics on CPU against the reference's own Python (SURVEY §4.2 'Unit')."""
import math

import pytest
import torch

from grace_b200 import compressor as C
from grace_b200.utils.philox import philox_uniform, philox_u32, random_subset, fnv1a, STREAM_UNIFORM, STREAM_SAMPLE

torch.manual_seed(1234)
X = torch.randn(96, 40)


def roundtrip(c, x=X, name="w"):
    payload, ctx = c.compress(x, name)
    return c.decompress(payload, ctx), payload


def test_philox_known_answer():
    # Random123 known-answer vectors for philox4x32-10
    from grace_b200.utils.philox import philox4x32
    z = philox4x32(torch.zeros(1, dtype=torch.int64), 0, 0, 0, 0, 0)
    assert [int(v) for v in z] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    f = philox4x32(torch.tensor([0xFFFFFFFF]), 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF)
    assert [int(v) for v in f] == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]


def test_philox_uniform_range_and_determinism():
    u = philox_uniform(10001, 7, fnv1a("conv1.weight"), 3)
    assert u.min() >= 0 and u.max() < 1 and abs(float(u.mean()) - 0.5) < 0.02
    assert torch.equal(u, philox_uniform(10001, 7, fnv1a("conv1.weight"), 3))
    assert not torch.equal(u, philox_uniform(10001, 7, fnv1a("conv1.weight"), 4))


@pytest.mark.parametrize("n,k", [(1, 1), (7, 3), (64, 64), (1000, 10), (4097, 500)])
def test_random_subset_is_without_replacement(n, k):
    idx = random_subset(n, k, 1, 2, 3)
    assert idx.numel() == k and idx.unique().numel() == k and idx.min() >= 0 and idx.max() < n


@pytest.mark.parametrize("ratio", [0.01, 0.1, 0.3])
def test_topk_matches_reference(reference, ratio):
    ref = reference("dist.compressor.topk").TopKCompressor(ratio)
    out, payload = roundtrip(C.TopKCompressor(ratio))
    exp, _ = roundtrip(ref)
    assert payload[0].numel() == max(1, int(X.numel() * ratio))
    assert torch.equal(out, exp)


def test_threshold_fp16_none_match_reference(reference):
    for ours, mod, cls, args in [(C.ThresholdCompressor(0.7), "threshold", "ThresholdCompressor", (0.7,)),
                                 (C.FP16Compressor(), "fp16", "FP16Compressor", ()),
                                 (C.NoneCompressor(), "none", "NoneCompressor", ())]:
        ref = getattr(reference("dist.compressor." + mod), cls)(*args)
        assert torch.equal(roundtrip(ours)[0], roundtrip(ref)[0])


def test_sign_family_matches_reference(reference):
    for ours, mod, cls, args in [(C.SignSGDCompressor(), "signsgd", "SignSGDCompressor", ()),
                                 (C.EFSignSGDCompressor(0.1), "efsignsgd", "EFSignSGDCompressor", (0.1,)),
                                 (C.SignumCompressor(0.9), "signum", "SignumCompressor", (0.9,))]:
        ref = getattr(reference("dist.compressor." + mod), cls)(*args)
        for _ in range(3):                               # momentum state evolves identically
            x = torch.randn(33, 7)
            a, b = roundtrip(ours, x)[0], roundtrip(ref, x)[0]
            assert torch.equal(a, b)
        parts = [torch.randn(5, 5).sign() for _ in range(4)]
        assert torch.equal(ours.aggregate(parts), ref.aggregate(parts))
        assert ours.average == ref.average and ours.tensors_size_are_same == ref.tensors_size_are_same


def test_onebit_follows_the_torch_flavour(reference):
    ref = reference("torch.compressor.onebit").OneBitCompressor()
    out, _ = roundtrip(C.OneBitCompressor())
    assert torch.allclose(out, roundtrip(ref)[0], rtol=1e-6, atol=1e-7)
    assert torch.allclose(roundtrip(C.OneBitCompressor(), X.abs())[0], X.abs().mean().expand_as(X))


def test_qsgd_same_arithmetic_as_reference_given_same_uniforms(reference, monkeypatch):
    for s in (8, 64, 256):
        ours = C.QSGDCompressor(s, seed=5)
        ref = reference("dist.compressor.qsgd").QSGDCompressor(s)
        u = philox_uniform(X.numel(), 5, fnv1a("w"), 0, STREAM_UNIFORM)
        monkeypatch.setattr(torch.Tensor, "uniform_", lambda self, *a, **k: self.copy_(u.view_as(self)))
        exp, rp = roundtrip(ref)
        monkeypatch.undo()
        out, p = roundtrip(ours)
        assert p[0].dtype == rp[0].dtype
        assert torch.allclose(out, exp, rtol=1e-6, atol=1e-7)


def test_qsgd_unbiased():
    c = C.QSGDCompressor(16)
    x = torch.randn(2000)
    acc = sum(roundtrip(c, x)[0] for _ in range(400)) / 400
    assert float((acc - x).abs().mean()) < 0.15 * float(x.abs().mean())


def test_terngrad_same_arithmetic_as_reference(reference, monkeypatch):
    ours = C.TernGradCompressor(seed=9)
    ref = reference("dist.compressor.terngrad").TernGradCompressor()
    u = philox_uniform(X.numel(), 9, fnv1a("w"), 0, STREAM_UNIFORM)
    from grace_b200.compressor.quant import terngrad_scalar
    scalar = float(terngrad_scalar(X.flatten()))
    monkeypatch.setattr(torch.Tensor, "uniform_", lambda self, *a, **k: self.copy_((u * scalar).view_as(self)))
    exp, rp = roundtrip(ref)
    monkeypatch.undo()
    out, p = roundtrip(ours)
    assert set(p[0].unique().tolist()) <= {-1, 0, 1}
    assert torch.allclose(p[1], rp[1], rtol=1e-6)
    assert float((out != exp).float().mean()) < 1e-3          # float rounding of u*scalar at the edge


def test_natural_is_unbiased_power_of_two():
    c = C.NaturalCompressor()
    x = torch.randn(4000) * 0.01
    out, p = roundtrip(c, x)
    assert p[0].dtype == torch.uint8
    nz = out[out != 0].abs()
    assert torch.equal(torch.log2(nz), torch.log2(nz).round())           # exact powers of two
    assert torch.all((out.abs() >= x.abs() / 2 - 1e-12) & (out.abs() <= x.abs() * 2 + 1e-12))
    acc = sum(roundtrip(c, x)[0] for _ in range(300)) / 300
    assert float((acc - x).abs().mean()) < 0.1 * float(x.abs().mean())


def test_natural_bit_arithmetic_matches_cupy_formulas():
    # restate natural.py:17-28 with numpy on the same random integers
    import numpy as np
    c = C.NaturalCompressor(seed=3)
    x = (torch.randn(513) * 3).float()
    payload, _ = c.compress(x, "n")
    r = (philox_u32(513, 3, fnv1a("n"), 0, STREAM_UNIFORM) >> 9).numpy().astype(np.int64)
    bits = x.numpy().view(np.int32).astype(np.int64) & 0xFFFFFFFF
    sign, exp, man = bits & 0x80000000, bits & 0x7F800000, bits & 0x007FFFFF
    exp = np.where(man > r, exp + 0x00800000, exp)
    exp = np.clip(exp, 0x09000000, 0x48800000) >> 23
    byte = ((sign >> 24) | (exp - 18)).astype(np.uint8)
    assert np.array_equal(byte, payload[0].numpy())


def test_u8bit_roundtrip_and_edges():
    c = C.U8bitCompressor()
    out, p = roundtrip(c)
    assert p[0].dtype == torch.int8 and float((out - X).norm() / X.norm()) < 0.05
    assert torch.equal(roundtrip(c, torch.zeros(10))[0], torch.zeros(10))
    big = torch.tensor([1.0, -1.0, 1e-9])
    assert torch.allclose(roundtrip(c, big)[0], torch.tensor([0.99570298, -0.99570298, 0.0]))


def test_sketch_bins_and_means():
    c = C.SketchCompressor(64)
    out, p = roundtrip(c)
    assert p[0].dtype == torch.uint8 and p[1].numel() == 64 and int(p[0].max()) == 63
    assert float((out - X).norm() / X.norm()) < 0.15
    counts = torch.bincount(p[0].flatten().long(), minlength=64)
    assert counts.max() - counts.min() <= 2                   # equal-mass buckets


def test_dgc_threshold_loop_matches_reference(reference, monkeypatch):
    ours = C.DgcCompressor(0.05, seed=11)
    ref = reference("dist.compressor.dgc").DgcCompressor(0.05)
    n = X.numel()
    u = philox_uniform(max(1, int(n * 0.01)), 11, fnv1a("w"), 0, STREAM_SAMPLE)
    monkeypatch.setattr(torch.Tensor, "uniform_", lambda self, lo, hi: self.copy_(u * hi))
    exp, rp = roundtrip(ref)
    monkeypatch.undo()
    out, p = roundtrip(ours)
    assert torch.equal(out, exp) and torch.equal(p[1], rp[1])


def test_adaq_payload_layout():
    c = C.AdaqCompressor(0.05)
    out, p = roundtrip(c)
    (wire,) = p
    assert wire.dtype == torch.int32
    n_plus = int(wire[2])
    plus, minus = wire[3:3 + n_plus].long(), wire[3 + n_plus:].long()
    flat = X.flatten()
    assert (flat[plus] > 0).all() and (flat[minus] < 0).all()
    assert torch.allclose(out.flatten()[plus], flat[plus].mean().expand(n_plus))
    assert int((out != 0).sum()) == wire.numel() - 3


def test_randomk_same_indices_for_same_name_and_step():
    a, b = C.RandomKCompressor(0.1, seed=1), C.RandomKCompressor(0.1, seed=1)
    state = torch.random.get_rng_state()
    pa, ca = a.compress(X, "layer")
    pb, cb = b.compress(X * 2, "layer")
    assert torch.equal(torch.random.get_rng_state(), state)     # global RNG untouched
    assert torch.equal(ca[0], cb[0]) and torch.equal(pb[0], 2 * pa[0])
    _, c2 = a.compress(X, "layer")
    assert not torch.equal(c2[0], ca[0])


def test_powersgd_exact_on_low_rank_and_eps_guard(reference):
    torch.manual_seed(0)
    m = torch.randn(40, 1) @ torch.randn(1, 24)
    c = C.PowerSGDCompressor(rank=4)
    out, p = roundtrip(c, m)
    assert p == [] and torch.isfinite(out).all()
    assert float((out - m).norm() / m.norm()) < 1e-4            # reference gives 0.56 here (no eps)
    vec = torch.randn(17)
    assert roundtrip(c, vec)[0] is vec
    ref_orth = reference("dist.compressor.powersgd").orthogonalize
    from grace_b200.compressor.lowrank import orthogonalize
    q = torch.randn(30, 3)
    a, b = q.clone(), q.clone()
    ref_orth(a)
    assert torch.allclose(orthogonalize(b), a, atol=1e-5)


def test_inceptionn_error_bound():
    c = C.INCEPTIONNCompressor(2e-10)
    x = torch.cat([torch.randn(500), torch.randn(500) * 0.05, torch.randn(500) * 1e-3])
    out, p = roundtrip(c, x)
    assert float((out - x).abs().max()) < 2 ** -5 + 1e-6
    big = x.abs() >= 1
    assert torch.equal(out[big], x[big])


@pytest.mark.parametrize("make", [lambda: C.TopKCompressor(0.1), lambda: C.QSGDCompressor(64), lambda: C.SignumCompressor(0.9),
                                  lambda: C.RandomKCompressor(0.2), lambda: C.PowerSGDCompressor(2, warm_start=True)])
def test_state_dict_roundtrip(make):
    a, b = make(), make()
    for _ in range(2):
        a.compress(X, "w")
    b.load_state_dict(a.state_dict())
    pa, ca = a.compress(X, "w")
    pb, cb = b.compress(X, "w")
    assert torch.allclose(a.decompress(pa, ca), b.decompress(pb, cb))


def test_signum_momentum_does_not_alias_the_callers_gradient():
    """The manual loop of the reference (core.py:202-208) overwrites the gradient in place after step();
    the first momentum must survive that."""
    c = C.SignumCompressor(0.9)
    g = torch.tensor([0.5, -0.25, 0.125])
    payload, ctx = c.compress(g, "w")
    g.copy_(c.decompress(payload, ctx))              # what `grad.copy_(grc.step(grad, name))` does
    g.zero_()                                        # optimizer.zero_grad(set_to_none=False)
    assert torch.equal(c.momentums["w"], torch.tensor([0.5, -0.25, 0.125]))
