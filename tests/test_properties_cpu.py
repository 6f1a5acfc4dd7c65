"""Property tests (hypothesis) for the codec layer -- SURVEY §4.2: shape/dtype preservation, the reference's
``k = max(1, int(numel * ratio))``, the residual identity, supports of the quantisers, Philox/permutation laws."""
import pytest
import torch

hypothesis = pytest.importorskip("hypothesis")
from hypothesis import given, settings, strategies as st   # noqa: E402

from grace_b200 import compressor as C                      # noqa: E402
from grace_b200 import memory as M                          # noqa: E402
from grace_b200.utils.philox import philox_uniform, STREAM_UNIFORM   # noqa: E402

shapes = st.lists(st.integers(1, 9), min_size=1, max_size=3)
seeds = st.integers(0, 2 ** 31 - 1)
ratios = st.sampled_from([0.001, 0.01, 0.05, 0.3, 1.0])


def _tensor(shape, seed, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


@settings(max_examples=40, deadline=None)
@given(shapes, seeds, ratios)
def test_topk_keeps_the_k_largest(shape, seed, ratio):
    x = _tensor(shape, seed)
    k = max(1, int(x.numel() * ratio))
    c = C.TopKCompressor(ratio)
    (vals, idx), ctx = c.compress(x, "w")
    assert vals.numel() == k == idx.numel() and idx.unique().numel() == k
    out = c.decompress((vals, idx), ctx)
    assert out.shape == x.shape and out.dtype == x.dtype
    kept = out.flatten() != 0
    if kept.any() and (~kept).any():
        assert x.flatten()[kept].abs().min() >= x.flatten()[~kept].abs().max()
    assert torch.equal(out.flatten()[idx], x.flatten()[idx])


@settings(max_examples=40, deadline=None)
@given(shapes, seeds, ratios, st.integers(0, 5))
def test_randomk_is_a_deterministic_subset_without_replacement(shape, seed, ratio, calls):
    x = _tensor(shape, seed)
    a, b = C.RandomKCompressor(ratio, seed=7), C.RandomKCompressor(ratio, seed=7)
    for _ in range(calls):                       # same call count on both "ranks"
        a.compress(x, "w"); b.compress(x, "w")
    (va,), ctxa = a.compress(x, "w")
    (vb,), ctxb = b.compress(x, "w")
    ia, ib = ctxa[0], ctxb[0]
    assert torch.equal(ia, ib) and torch.equal(va, vb)
    assert ia.unique().numel() == ia.numel() == max(1, int(x.numel() * ratio))
    assert int(ia.min()) >= 0 and int(ia.max()) < x.numel()


@settings(max_examples=30, deadline=None)
@given(shapes, seeds, st.sampled_from([4, 64, 127, 128, 256]))
def test_qsgd_levels_are_integers_in_range(shape, seed, s):
    x = _tensor(shape, seed)
    c = C.QSGDCompressor(s, seed=3)
    (q, norm), ctx = c.compress(x, "w")
    assert float(q.float().abs().max()) <= s
    out = c.decompress((q, norm), ctx)
    assert out.shape == x.shape
    assert (out.sign() * x.sign() >= 0).all()                      # never flips a sign
    assert float((out - x).abs().max()) <= float(norm) / s * (1 + 1e-3) + 1e-12   # at most one level away


@settings(max_examples=30, deadline=None)
@given(shapes, seeds)
def test_terngrad_support(shape, seed):
    x = _tensor(shape, seed)
    c = C.TernGradCompressor(seed=5)
    (t, scalar), ctx = c.compress(x, "w")
    assert set(t.flatten().tolist()) <= {-1, 0, 1}
    out = c.decompress((t, scalar), ctx)
    vals = set(round(v, 6) for v in (out / (scalar + 1e-30)).flatten().tolist())
    assert vals <= {-1.0, 0.0, 1.0}


@settings(max_examples=30, deadline=None)
@given(shapes, seeds)
def test_natural_outputs_signed_powers_of_two(shape, seed):
    x = _tensor(shape, seed)
    c = C.NaturalCompressor(seed=11)
    payload, ctx = c.compress(x, "w")
    out = c.decompress(payload, ctx).flatten()
    nz = out[out != 0]
    if nz.numel():
        m, _ = torch.frexp(nz.abs())
        assert torch.all(m == 0.5)                                  # exact powers of two
        ratio = nz.abs() / x.flatten()[out != 0].abs()
        assert float(ratio.min()) > 0.5 - 1e-6 and float(ratio.max()) <= 2.0 + 1e-6
    assert (out.sign() * x.flatten().sign() >= 0).all()


@settings(max_examples=25, deadline=None)
@given(seeds, st.sampled_from(["topk", "qsgd", "signsgd", "natural", "threshold", "u8bit", "adaq"]), st.integers(1, 4))
def test_residual_identity_over_several_steps(seed, codec, steps):
    make = {"topk": lambda: C.TopKCompressor(0.1), "qsgd": lambda: C.QSGDCompressor(64), "signsgd": C.SignSGDCompressor,
            "natural": C.NaturalCompressor, "threshold": lambda: C.ThresholdCompressor(0.7), "u8bit": C.U8bitCompressor,
            "adaq": lambda: C.AdaqCompressor(0.1)}[codec]
    comp, mem = make(), M.ResidualMemory(beta=0.9, gamma=1.1)
    g = torch.Generator().manual_seed(seed)
    for _ in range(steps):
        x = mem.compensate(torch.randn(64, generator=g), "w")
        payload, ctx = comp.compress(x, "w")
        mem.update(x, "w", comp, payload, ctx)
        assert torch.allclose(comp.decompress(payload, ctx).flatten() + mem.residuals["w"].flatten(), x.flatten(), atol=1e-6, rtol=0)


@settings(max_examples=40, deadline=None)
@given(st.integers(1, 5000), seeds, st.integers(0, 2 ** 32 - 1), st.integers(0, 1000))
def test_philox_uniform_is_a_pure_function_of_its_counter(n, seed, name_hash, step):
    a = philox_uniform(n, seed, name_hash, step, STREAM_UNIFORM)
    b = philox_uniform(n + 13, seed, name_hash, step, STREAM_UNIFORM)
    assert torch.equal(a, b[:n])                                     # prefix-stable: element i depends only on i
    assert float(a.min()) >= 0.0 and float(a.max()) < 1.0
    if n >= 8:
        c = philox_uniform(n, seed, name_hash, step + 1, STREAM_UNIFORM)
        assert not torch.equal(a, c)
