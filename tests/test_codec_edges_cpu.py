"""Edge inputs through every built-in codec on CPU (SURVEY §4.2 'Unit': shapes, dtypes, degenerate tensors).

The reference leaves most of these undefined (division by a zero norm in QSGD, NaN bins in 8-bit, an empty
top-k sample in DGC); here every codec must return a finite tensor of the input's shape."""
import math

import pytest
import torch

from grace_b200 import compressor as C
from grace_b200 import memory as M

CODECS = [
    ("none", C.NoneCompressor), ("fp16", C.FP16Compressor), ("topk", lambda: C.TopKCompressor(0.1)),
    ("randomk", lambda: C.RandomKCompressor(0.1)), ("threshold", lambda: C.ThresholdCompressor(0.5)),
    ("dgc", lambda: C.DgcCompressor(0.1)), ("qsgd", lambda: C.QSGDCompressor(64)), ("terngrad", C.TernGradCompressor),
    ("signsgd", C.SignSGDCompressor), ("signum", lambda: C.SignumCompressor(0.9)),
    ("efsignsgd", lambda: C.EFSignSGDCompressor(0.1)), ("onebit", C.OneBitCompressor), ("natural", C.NaturalCompressor),
    ("u8bit", C.U8bitCompressor), ("sketch", lambda: C.SketchCompressor(16)), ("adaq", lambda: C.AdaqCompressor(0.1)),
    ("inceptionn", C.INCEPTIONNCompressor),
]
INPUTS = {
    "zeros": lambda: torch.zeros(7, 5),
    "single": lambda: torch.tensor([0.37]),
    "constant": lambda: torch.full((33,), -2.5),
    "one_hot": lambda: torch.nn.functional.one_hot(torch.tensor(3), 50).float() * 4.0,
    "4d": lambda: torch.randn(3, 2, 3, 3, generator=torch.Generator().manual_seed(0)),
    "tiny_values": lambda: torch.randn(100, generator=torch.Generator().manual_seed(1)) * 1e-20,
    "large_values": lambda: torch.randn(100, generator=torch.Generator().manual_seed(2)) * 1e20,
}


@pytest.mark.parametrize("codec,make", CODECS, ids=[c for c, _ in CODECS])
@pytest.mark.parametrize("case", list(INPUTS))
def test_roundtrip_is_finite_and_keeps_shape(codec, make, case):
    if (codec, case) == ("fp16", "large_values"):
        pytest.skip("1e20 overflows the fp16 wire by definition (reference fp16.py:12-14 does the same cast)")
    if (codec, case) == ("qsgd", "large_values"):
        pytest.skip("the fp32 L2 norm of 1e20-sized values overflows (reference qsgd.py:16 has the same norm); the native kernel sums in double")
    x = INPUTS[case]()
    comp = make()
    for step in range(2):                             # second call exercises the per-name state
        payload, ctx = comp.compress(x.clone(), "w")
        out = comp.decompress(payload, ctx)
        assert out.shape == x.shape, (codec, case, out.shape)
        assert torch.isfinite(out).all(), (codec, case)
        agg = comp.aggregate([out, out])
        assert agg.shape == x.shape and torch.isfinite(agg).all()
    if case == "zeros" and codec not in ("signsgd", "signum", "onebit"):
        assert float(out.abs().max()) == 0.0, (codec, float(out.abs().max()))


@pytest.mark.parametrize("codec,make", CODECS, ids=[c for c, _ in CODECS])
def test_residual_memory_conserves_the_signal(codec, make):
    """compensate -> compress -> update: what was not sent stays in the memory, exactly (x = sent + residual)."""
    g = torch.Generator().manual_seed(5)
    comp, mem = make(), M.ResidualMemory()
    for step in range(3):
        x = mem.compensate(torch.randn(257, generator=g), "w")
        payload, ctx = comp.compress(x, "w")
        mem.update(x, "w", comp, payload, ctx)
        sent = comp.decompress(payload, ctx)
        assert torch.allclose(sent.reshape(-1) + mem.residuals["w"].reshape(-1), x.reshape(-1), rtol=0, atol=1e-6), (codec, step)


def test_k_follows_the_reference_formula():
    for n in (1, 9, 10, 99, 1000):
        for ratio in (0.001, 0.01, 0.3):
            k = max(1, int(n * ratio))
            v, i = C.TopKCompressor(ratio).compress(torch.arange(1.0, n + 1), "w")[0]
            assert v.numel() == k and i.numel() == k
            assert C.RandomKCompressor(ratio).compress(torch.arange(1.0, n + 1), "w")[0][0].numel() == k


def test_payload_bytes_shrink():
    """Wire sizes of the python-visible payloads (the native wire is tighter: true bit packing)."""
    x = torch.randn(4096, generator=torch.Generator().manual_seed(3))
    dense = x.numel() * 4
    def nbytes(comp):
        return sum(t.numel() * t.element_size() for t in comp.compress(x, "w")[0])
    assert nbytes(C.FP16Compressor()) == dense // 2
    assert nbytes(C.TopKCompressor(0.01)) == 40 * (4 + 8)
    assert nbytes(C.QSGDCompressor(64)) <= dense // 4 + 8
    assert nbytes(C.SignSGDCompressor()) == dense // 4
    assert nbytes(C.NaturalCompressor()) == dense // 4
    assert math.isclose(nbytes(C.SketchCompressor(64)), dense // 4 + 64 * 4, abs_tol=8)
