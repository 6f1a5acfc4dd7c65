"""INCEPTIONN codec (TensorFlow-only in the reference:
``grace_dl/tensorflow/compressor/inceptionn.py:8-188``; optional).

Each value is sent with 32, 16, 8 or 0 bits depending on its binary exponent relative to
an ``error_bound``; a 2-bit tag per element (packed four to a byte, like
``tensorflow/compressor/packing.py:4-30``) says which.  The 16- and 8-bit forms are
sign + truncated fixed point of ``|x| < 1`` (scale 2**14 and 2**6), which is what the
reference's shift-and-marker arithmetic produces; here it is written as plain
arithmetic on the magnitudes.
"""
from __future__ import annotations

import math

import torch

from grace_b200.core import Compressor


def pack_tags(tags):
    """2-bit tags -> bytes: element ``i`` of quarter ``j`` lands in bits ``2j`` of byte ``i``."""
    n = tags.numel()
    pad = 4 - n % 4
    tags = torch.cat([tags.to(torch.int32), torch.arange(pad, dtype=torch.int32, device=tags.device)])
    a, b, c, d = tags.view(4, -1)
    return (a + b * 4 + c * 16 + d * 64).to(torch.uint8)


def unpack_tags(packed, n):
    v = packed.to(torch.int32)
    return torch.cat([v % 4, (v // 4) % 4, (v // 16) % 4, (v // 64) % 4])[:n]


class INCEPTIONNCompressor(Compressor):
    def __init__(self, error_bound=2e-10):
        super().__init__(tensors_size_are_same=False)
        self.error_bound = error_bound
        self.low_exp = 127 + int(math.log(error_bound / 2, 10))
        self.mid_exp = self.low_exp + math.ceil((127 - self.low_exp) / 2)

    def compress(self, tensor, name):
        flat = tensor.flatten().float().contiguous()
        expo = (flat.view(torch.int32).to(torch.int64) >> 23) & 0xFF
        m32 = expo >= 127
        m16 = (expo >= self.mid_exp) & ~m32
        m8 = (expo >= self.low_exp) & (expo < self.mid_exp)
        neg = flat < 0
        mag = flat.abs()
        v32 = flat[m32]
        c16 = (mag[m16] * 16384.0).floor().to(torch.int32) | (neg[m16].to(torch.int32) << 15)
        c8 = (mag[m8] * 64.0).floor().to(torch.int32) | (neg[m8].to(torch.int32) << 7)
        tags = m8.to(torch.int32) + 2 * m16.to(torch.int32) + 3 * m32.to(torch.int32)
        return [v32, c16.to(torch.int16), c8.to(torch.uint8), pack_tags(tags)], tensor.size()

    def decompress(self, tensors, ctx):
        v32, c16, c8, packed = tensors
        numel = 1
        for d in ctx:
            numel *= int(d)
        tags = unpack_tags(packed, numel)
        out = torch.zeros(numel, dtype=torch.float32, device=packed.device)
        out[tags == 3] = v32
        w16 = c16.to(torch.int32) & 0xFFFF
        out[tags == 2] = (w16 & 0x7FFF).to(torch.float32) / 16384.0 * (1 - 2 * (w16 >> 15)).to(torch.float32)
        w8 = c8.to(torch.int32)
        out[tags == 1] = (w8 & 0x7F).to(torch.float32) / 64.0 * (1 - 2 * (w8 >> 7)).to(torch.float32)
        return out.view(ctx)
