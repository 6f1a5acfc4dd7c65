"""Counter-based randomness shared bit-for-bit between the torch oracle and the CUDA kernels.

The reference draws from the global torch / CuPy generators (``grace_dl/dist/compressor/
qsgd.py:20``, ``terngrad.py:20``, ``natural.py:21``) and RandomK even *reseeds* the
global generator on every call (``randomk.py:26-28``).  Here every
stochastic codec consumes a private Philox4x32-10 stream keyed by

    key     = (seed, fnv1a(name))
    counter = (element_index // 4, step, stream_id, 0)        lane = element_index % 4

so (a) all ranks can derive identical draws from (name, step) without touching global
state, and (b) the fused kernels in ``csrc/`` reproduce exactly the numbers the torch
restatement sees, which makes the stochastic kernels testable bit-for-bit.

The same constants and bit manipulations are implemented in ``csrc/philox.cuh``.
"""
from __future__ import annotations

import torch

_M0 = 0xD2511F53
_M1 = 0xCD9E8D57
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = 0xFFFFFFFF

STREAM_UNIFORM = 0   # QSGD / TernGrad / Natural per-element draws
STREAM_SAMPLE = 1    # DGC / AdaQ sampling positions
STREAM_NORMAL = 2    # PowerSGD Q initialisation
STREAM_PERMUTE = 3   # RandomK round keys


def fnv1a(name: str) -> int:
    """32-bit FNV-1a of the UTF-8 bytes of ``name`` (rank-invariant name id)."""
    h = 0x811C9DC5
    for b in name.encode("utf8"):
        h = ((h ^ b) * 0x01000193) & _MASK
    return h


def _mulhilo(a: int, b: torch.Tensor):
    # b holds uint32 values in int64; the product wraps mod 2^64 which is what we want
    lo_part = (b & 0xFFFF) * a
    hi_part = (b >> 16) * a
    # 64-bit product = hi_part * 2^16 + lo_part, computed without overflowing int64's sign
    lo = (lo_part + ((hi_part & 0xFFFF) << 16))
    carry = lo >> 32
    hi = ((hi_part >> 16) + carry) & _MASK
    return hi, lo & _MASK


def philox4x32(c0: torch.Tensor, c1: int, c2: int, c3: int, k0: int, k1: int):
    """Vectorised Philox4x32-10.  ``c0`` is an int64 tensor of uint32 counters; the
    other counter words and the key are python ints.  Returns four int64 tensors
    holding uint32 outputs."""
    c0 = c0.to(torch.int64) & _MASK
    c1 = torch.full_like(c0, c1 & _MASK)
    c2 = torch.full_like(c0, c2 & _MASK)
    c3 = torch.full_like(c0, c3 & _MASK)
    k0 &= _MASK
    k1 &= _MASK
    for _ in range(10):
        hi0, lo0 = _mulhilo(_M0, c0)
        hi1, lo1 = _mulhilo(_M1, c2)
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & _MASK, lo1, (hi0 ^ c3 ^ k1) & _MASK, lo0
        k0 = (k0 + _W0) & _MASK
        k1 = (k1 + _W1) & _MASK
    return c0, c1, c2, c3


def philox_u32(n: int, seed: int, name_hash: int, step: int, stream: int, device="cpu") -> torch.Tensor:
    """``n`` uint32 draws (as int64) for element indices ``0..n-1``."""
    n4 = (n + 3) // 4
    ctr = torch.arange(n4, dtype=torch.int64, device=device)
    r = philox4x32(ctr, step, stream, 0, seed, name_hash)
    return torch.stack(r, dim=1).reshape(-1)[:n]


def philox_uniform(n: int, seed: int, name_hash: int, step: int, stream: int = STREAM_UNIFORM,
                   device="cpu") -> torch.Tensor:
    """float32 uniforms in [0, 1) with 24 bits of resolution: ``(x >> 8) * 2**-24``."""
    x = philox_u32(n, seed, name_hash, step, stream, device)
    return (x >> 8).to(torch.float32) * (1.0 / 16777216.0)


def philox_normal(n: int, seed: int, name_hash: int, step: int, stream: int = STREAM_NORMAL,
                  device="cpu") -> torch.Tensor:
    """One standard normal per element by Box-Muller on lanes 0/1 of counter = element."""
    ctr = torch.arange(n, dtype=torch.int64, device=device)
    x0, x1, _, _ = philox4x32(ctr, step, stream, 0, seed, name_hash)
    u1 = ((x0 >> 8) + 1).to(torch.float32) * (1.0 / 16777216.0)      # (0, 1]
    u2 = (x1 >> 8).to(torch.float32) * (1.0 / 16777216.0)            # [0, 1)
    return torch.sqrt(-2.0 * torch.log(u1)) * torch.cos(6.283185307179586 * u2)


def _lowbias32(x: torch.Tensor) -> torch.Tensor:
    x = x & _MASK
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & _MASK
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & _MASK
    x = x ^ (x >> 16)
    return x


def feistel_bits(n: int) -> int:
    """Half-width of the Feistel domain covering ``[0, n)`` (domain = 4**half >= n)."""
    b = max(2, (max(n, 2) - 1).bit_length())
    if b & 1:
        b += 1
    return b // 2


def random_subset(n: int, k: int, seed: int, name_hash: int, step: int, device="cpu") -> torch.Tensor:
    """First ``k`` images of a keyed bijection of ``[0, n)`` -- a without-replacement
    sample that every rank derives identically from (name, step).  Four-round Feistel
    network on ``2*half`` bits with cycle walking; round keys come from Philox."""
    half = feistel_bits(n)
    mask = (1 << half) - 1
    keys = philox4x32(torch.zeros(1, dtype=torch.int64), step, STREAM_PERMUTE, 0, seed, name_hash)
    keys = [int(x.item()) for x in keys]
    x = torch.arange(k, dtype=torch.int64, device=device)

    def rounds(v):
        left, right = (v >> half) & mask, v & mask
        for rk in keys:
            left, right = right, left ^ (_lowbias32(right ^ rk) & mask)
        return (left << half) | right

    out = rounds(x)
    for _ in range(256):
        bad = out >= n
        if not bool(bad.any()):
            break
        out = torch.where(bad, rounds(out), out)
    return out
