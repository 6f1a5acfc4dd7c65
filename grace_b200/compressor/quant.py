"""Quantising codecs: QSGD, TernGrad, Natural, 8-bit (U8bit), SketchML.

Torch restatements defining the numerics of the fused K2 kernels (``csrc/k2_quant.cu``).
Reference files: ``grace_dl/dist/compressor/qsgd.py:6-38``, ``terngrad.py:6-32``,
``natural.py:9-40`` (CuPy) and, TensorFlow only, ``grace_dl/tensorflow/compressor/
u8bit.py:6-110`` and ``sketch.py:6-39``.

The Python-visible payload keeps the reference's one-value-per-element dtypes; the fused
path bit-packs the same codes on the wire.
"""
from __future__ import annotations

import torch

from grace_b200.core import Compressor
from grace_b200.compressor._common import StochasticCompressor, DEFAULT_SEED
from grace_b200.utils.philox import philox_uniform, philox_u32, STREAM_UNIFORM


class QSGDCompressor(StochasticCompressor):
    """Stochastic uniform quantisation of ``|x|/||x||`` onto ``quantum_num`` levels."""

    def __init__(self, quantum_num, seed=DEFAULT_SEED):
        super().__init__(seed=seed)
        self.quantum_num = quantum_num

    def compress(self, tensor, name):
        flat = tensor.flatten().float()
        nh, step = self._tick(name)
        u = philox_uniform(flat.numel(), self.seed, nh, step, STREAM_UNIFORM, device=flat.device)
        norm = flat.norm()
        scale = torch.where(norm > 0, self.quantum_num / norm, torch.zeros_like(norm))
        level = scale * flat.abs()
        low = level.floor()
        level_q = low + (u < (level - low)).to(torch.float32)
        wire_dtype = torch.int8 if self.quantum_num < 128 else torch.float16
        q = (level_q * flat.sign()).to(torch.int16).to(wire_dtype)
        return [q, norm], tensor.size()

    def decompress(self, tensors, ctx):
        q, norm = tensors
        return (norm / self.quantum_num * q.to(torch.float32)).view(ctx)


def terngrad_scalar(flat):
    """``min(max|x|, 2.5 sigma)`` -- what ``clamp(x, +-c).abs().max()`` evaluates to
    (``terngrad.py:12-17``), written so that a kernel needs one reduction pass."""
    sigma = torch.sqrt(torch.mean((flat - flat.mean()) ** 2))
    return torch.minimum(flat.abs().max(), 2.5 * sigma)


class TernGradCompressor(StochasticCompressor):
    """Ternary {-1, 0, +1} codes times one scalar, with 2.5-sigma clipping."""

    def __init__(self, seed=DEFAULT_SEED):
        super().__init__(seed=seed)

    def compress(self, tensor, name):
        flat = tensor.flatten().float()
        nh, step = self._tick(name)
        u = philox_uniform(flat.numel(), self.seed, nh, step, STREAM_UNIFORM, device=flat.device)
        scalar = terngrad_scalar(flat)
        mag = torch.minimum(flat.abs(), scalar)
        keep = (u * scalar) < mag
        code = torch.where(keep, flat.sign(), torch.zeros_like(flat)).to(torch.int8)
        return [code, scalar.reshape(1)], tensor.size()

    def decompress(self, tensors, ctx):
        code, scalar = tensors
        return (code.to(torch.float32) * scalar).view(ctx)


class NaturalCompressor(StochasticCompressor):
    """Natural compression: stochastic rounding to a power of two, one byte per value
    (sign bit + 7-bit exponent offset).  Pure integer arithmetic on the float bits."""

    def __init__(self, seed=DEFAULT_SEED):
        super().__init__(seed=seed)

    def compress(self, tensor, name):
        flat = tensor.flatten().float().contiguous()
        nh, step = self._tick(name)
        r = philox_u32(flat.numel(), self.seed, nh, step, STREAM_UNIFORM, device=flat.device) >> 9
        bits = flat.view(torch.int32).to(torch.int64) & 0xFFFFFFFF
        sign = bits >> 31
        expo = (bits >> 23) & 0xFF
        mant = bits & 0x7FFFFF
        expo = (expo + (mant > r).to(torch.int64)).clamp_(18, 145)
        return [((sign << 7) | (expo - 18)).to(torch.uint8)], tensor.size()

    def decompress(self, tensors, ctx):
        (byte,) = tensors
        b = byte.to(torch.int64)
        e = b & 0x7F
        mag = ((e + 18) << 23).to(torch.int32).view(torch.float32)
        val = torch.where(b > 127, -mag, mag)
        return torch.where(e >= 1, val, torch.zeros_like(val)).view(ctx)


# Dettmers' dynamic 8-bit magnitude table (arXiv:1511.04561), 128 ascending entries.
# These are the data values of ``tensorflow/compressor/u8bit.py:12-47``.
_U8_TABLE = (
    1.5000001e-06, 2.75e-06, 7.2499997e-06, 1.8750001e-05, 3.625e-05, 5.8749996e-05, 8.6249995e-05,
    1.4375e-04, 2.3125e-04, 3.1875001e-04, 4.0625001e-04, 5.1874999e-04, 6.5624999e-04, 7.9374999e-04,
    9.3124999e-04, 1.21875e-03, 1.65625e-03, 2.0937501e-03, 2.53125e-03, 2.96875e-03, 3.4062499e-03,
    3.8437501e-03, 4.2812498e-03, 4.84375e-03, 5.53125e-03, 6.21875e-03, 6.90625e-03, 7.59375e-03,
    8.2812496e-03, 8.96875e-03, 9.6562495e-03, 1.109375e-02, 1.328125e-02, 1.546875e-02, 1.765625e-02,
    1.984375e-02, 2.2031249e-02, 2.4218749e-02, 2.6406251e-02, 2.8593751e-02, 3.078125e-02, 3.2968748e-02,
    3.515625e-02, 3.7343752e-02, 3.953125e-02, 4.1718751e-02, 4.3906249e-02, 4.671875e-02, 5.0156251e-02,
    5.3593751e-02, 5.7031251e-02, 6.0468748e-02, 6.3906237e-02, 6.7343749e-02, 7.0781253e-02,
    7.4218743e-02, 7.7656247e-02, 8.1093743e-02, 8.453124e-02, 8.7968737e-02, 9.1406241e-02,
    9.4843738e-02, 9.8281242e-02, 1.0546875e-01, 1.1640625e-01, 1.2734374e-01, 1.3828126e-01,
    1.4921875e-01, 1.6015625e-01, 1.7109375e-01, 1.8203124e-01, 1.9296876e-01, 2.0390625e-01,
    2.1484375e-01, 2.2578125e-01, 2.3671874e-01, 2.4765626e-01, 2.5859374e-01, 2.6953125e-01,
    2.8046876e-01, 2.9140624e-01, 3.0234376e-01, 3.1328124e-01, 3.2421875e-01, 3.3515626e-01,
    3.4609374e-01, 3.5703126e-01, 3.6796874e-01, 3.7890625e-01, 3.8984376e-01, 4.0078124e-01,
    4.1171876e-01, 4.2265624e-01, 4.3359375e-01, 4.4453126e-01, 4.5859376e-01, 4.7578123e-01,
    4.9296874e-01, 5.1015621e-01, 5.2734375e-01, 5.4453123e-01, 5.617187e-01, 5.7890624e-01,
    5.9609371e-01, 6.1328125e-01, 6.3046873e-01, 6.476562e-01, 6.6484374e-01, 6.8203121e-01,
    6.9921869e-01, 7.1640623e-01, 7.335937e-01, 7.5078118e-01, 7.6796871e-01, 7.8515619e-01,
    8.0234367e-01, 8.195312e-01, 8.3671868e-01, 8.5390615e-01, 8.7109369e-01, 8.8828117e-01,
    9.0546864e-01, 9.2265618e-01, 9.3984365e-01, 9.5703113e-01, 9.7421867e-01, 9.9140614e-01,
    9.9570298e-01,
)


def u8_table(device="cpu"):
    return torch.tensor(_U8_TABLE, dtype=torch.float32, device=device)


class U8bitCompressor(Compressor):
    """Dynamic-tree 8-bit codes: ``code = sign * bin(|x| / max|x|)`` over a fixed
    non-uniform 128-entry table.  Out-of-table magnitudes are clamped to the first/last
    bin (the reference leaves them undefined)."""

    def compress(self, tensor, name):
        flat = tensor.flatten().float()
        scaler = flat.abs().max()
        inv = torch.where(scaler > 0, 1.0 / scaler, torch.zeros_like(scaler))
        y = (flat.abs() * inv).contiguous()
        bins = (torch.searchsorted(u8_table(flat.device), y, right=True) - 1).clamp_(0, 127)
        code = (bins * flat.sign().to(torch.int64)).to(torch.int8)
        return [code, scaler.reshape(1)], tensor.size()

    def decompress(self, tensors, ctx):
        code, scaler = tensors
        c = code.to(torch.int64)
        mag = u8_table(code.device)[c.abs().clamp_(max=127)]
        return (mag * scaler * c.sign().to(torch.float32)).view(ctx)


def linear_quantile_edges(flat, q):
    """``q + 1`` cut points min..max with linear interpolation between order statistics
    (the 'linear' mode of ``tfp.stats.quantiles`` used at ``sketch.py:22``)."""
    srt = torch.sort(flat)[0]
    n = srt.numel()
    pos = torch.arange(q + 1, dtype=torch.float64, device=flat.device) * ((n - 1) / q)
    lo = pos.floor().to(torch.int64)
    hi = pos.ceil().to(torch.int64).clamp_(max=n - 1)
    frac = (pos - lo.to(torch.float64)).to(torch.float32)
    return srt[lo] + (srt[hi] - srt[lo]) * frac


class SketchCompressor(Compressor):
    """SketchML: quantile bucketing; each value is replaced by its bucket's mean."""

    def __init__(self, quantiles=64):
        super().__init__()
        self.quantiles = quantiles

    def compress(self, tensor, name):
        flat = tensor.flatten().float().contiguous()
        q = self.quantiles
        edges = linear_quantile_edges(flat, q)
        bins = (torch.searchsorted(edges.contiguous(), flat, right=True) - 1).clamp_(0, q - 1)
        sums = torch.zeros(q, dtype=torch.float32, device=flat.device).index_add_(0, bins, flat)
        cnts = torch.zeros(q, dtype=torch.float32, device=flat.device).index_add_(0, bins, torch.ones_like(flat))
        means = torch.where(cnts > 0, sums / cnts.clamp(min=1), torch.zeros_like(sums))
        return [bins.to(torch.uint8 if q < 256 else torch.int16), means], tensor.size()

    def decompress(self, tensors, ctx):
        bins, means = tensors
        return means[bins.to(torch.int64)].view(ctx)
