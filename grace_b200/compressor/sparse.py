"""Sparsifying codecs: TopK, RandomK, Threshold, DGC, AdaQ.

Torch restatements that define the numerics the fused K1 kernels (``csrc/k1_sparse.cu``)
must reproduce.  Reference files: ``grace_dl/dist/compressor/topk.py:6-36``,
``randomk.py:6-40``, ``threshold.py:6-27``, ``dgc.py:6-50`` and (TensorFlow only)
``grace_dl/tensorflow/compressor/adaq.py:6-93``.

Deliberate deviations: RandomK uses a private keyed permutation instead of
reseeding the global generator (item 8); DGC/AdaQ sample positions come from the same
Philox stream the kernels use; DGC follows the torch constants, AdaQ the TF ones (item 16).
"""
from __future__ import annotations

import math

import torch

from grace_b200.core import Compressor
from grace_b200.compressor._common import StochasticCompressor, scatter_dense, topk_count, DEFAULT_SEED
from grace_b200.utils.philox import philox_uniform, random_subset, STREAM_SAMPLE


class TopKCompressor(Compressor):
    """Keep the ``k = max(1, int(numel*ratio))`` largest-magnitude entries."""

    def __init__(self, compress_ratio):
        super().__init__()
        self.compress_ratio = compress_ratio

    def compress(self, tensor, name):
        flat = tensor.flatten()
        k = topk_count(flat.numel(), self.compress_ratio)
        _, indices = torch.topk(flat.abs(), k, sorted=False)
        return [flat[indices], indices], (flat.numel(), tensor.size())

    def decompress(self, tensors, ctx):
        numel, shape = ctx
        values, indices = tensors
        return scatter_dense(values, indices, numel).view(shape)


class RandomKCompressor(StochasticCompressor):
    """Send ``k`` values at positions every rank derives from ``(name, call#)``; the
    positions themselves never travel.  Linear, hence all-reduce compatible."""

    def __init__(self, compress_ratio, seed=DEFAULT_SEED):
        super().__init__(seed=seed)
        self.compress_ratio = compress_ratio

    @property
    def global_step(self):  # kept for parity with the reference attribute
        return sum(self.steps.values())

    def compress(self, tensor, name):
        flat = tensor.flatten()
        k = topk_count(flat.numel(), self.compress_ratio)
        nh, step = self._tick(name)
        indices = random_subset(flat.numel(), k, self.seed, nh, step, device=flat.device)
        return [flat[indices]], (indices, flat.numel(), tensor.size())

    def decompress(self, tensors, ctx):
        indices, numel, shape = ctx
        (values,) = tensors
        return scatter_dense(values, indices, numel).view(shape)


class ThresholdCompressor(Compressor):
    """Send every entry with ``|x| > threshold``; payload size varies per rank."""

    def __init__(self, threshold):
        super().__init__(tensors_size_are_same=False)
        self.threshold = threshold

    def compress(self, tensor, name):
        flat = tensor.flatten()
        (indices,) = torch.where(flat.abs() > self.threshold)
        return [flat[indices], indices], (tensor.size(), flat.numel())

    def decompress(self, tensors, ctx):
        shape, numel = ctx
        values, indices = tensors
        return scatter_dense(values, indices, numel).view(shape)


def dgc_threshold(flat, ratio, sample_u):
    """Sampled top-k threshold followed by the multiplicative refinement loop of
    ``grace_dl/dist/compressor/dgc.py:17-36``.  ``sample_u`` are uniforms in [0,1)."""
    numel = flat.numel()
    absx = flat.abs()
    pos = (sample_u * numel).to(torch.int64).clamp_(max=numel - 1)
    k = max(1, int(numel * ratio * 0.01))
    k = min(k, pos.numel())
    thr = torch.topk(absx[pos], k)[0].min()
    mask = absx >= thr
    selected = int(mask.sum())
    for _ in range(10):
        if selected > 1.3 * numel * ratio:
            thr = 1.3 * thr
        elif selected < 0.7 * numel * ratio:
            thr = 0.7 * thr
        else:
            break
        mask = absx >= thr
        selected = int(mask.sum())
    return thr, mask


class DgcCompressor(StochasticCompressor):
    """Deep Gradient Compression selection; ``ctx`` carries the mask that
    ``DgcMemory.update`` consumes."""

    def __init__(self, compress_ratio, seed=DEFAULT_SEED):
        super().__init__(tensors_size_are_same=False, seed=seed)
        self.compress_ratio = compress_ratio

    def compress(self, tensor, name):
        shape = tensor.size()
        flat = tensor.flatten()
        numel = flat.numel()
        nh, step = self._tick(name)
        ns = max(1, int(numel * 0.01))
        u = philox_uniform(ns, self.seed, nh, step, STREAM_SAMPLE, device=flat.device)
        _, mask = dgc_threshold(flat, self.compress_ratio, u)
        (indices,) = torch.where(mask)
        return [flat[indices], indices], (shape, mask, numel)

    def decompress(self, tensors, ctx):
        values, indices = tensors
        shape, _, numel = ctx
        return scatter_dense(values, indices, numel).view(shape)


def _adaq_part(flat, part_mask, ratio, u):
    """One signed half of AdaQ (``tensorflow/compressor/adaq.py:16-55``)."""
    count = int(part_mask.sum())
    if count == 0:
        empty = torch.empty(0, dtype=torch.int64, device=flat.device)
        return empty, torch.zeros((), dtype=flat.dtype, device=flat.device)
    vals = flat[part_mask]
    ns = int(math.ceil(count * 0.01))
    pos = (u[:ns] * count).to(torch.int64).clamp_(max=count - 1)
    k = max(1, min(ns, int(math.ceil(count * 0.01 * ratio))))
    thr = torch.topk(vals[pos].abs(), k)[0].min()
    masked = torch.where(part_mask, flat, torch.zeros_like(flat)).abs()
    target = math.ceil(count * ratio)
    mask = masked > thr
    selected = int(mask.sum())
    for _ in range(20):
        if not (selected > 1.25 * target or selected < 0.8 * target):
            break
        thr = 1.25 * thr if selected > 1.25 * target else 0.9 * thr
        mask = masked >= thr
        selected = int(mask.sum())
    if selected < 1:
        thr = 0.8 * thr
    mask = masked > thr
    (indices,) = torch.where(mask)
    mean = flat[indices].mean() if indices.numel() else torch.zeros((), dtype=flat.dtype, device=flat.device)
    return indices, mean


class AdaqCompressor(StochasticCompressor):
    """AdaComp-style: per sign, keep only the positions above a sampled threshold and a
    single mean.  Wire = one int32 vector ``[mean+, mean-, n+, idx+..., idx-...]``."""

    def __init__(self, compress_ratio, seed=DEFAULT_SEED):
        super().__init__(tensors_size_are_same=False, seed=seed)
        self.compress_ratio = compress_ratio

    def compress(self, tensor, name):
        flat = tensor.flatten().float()
        numel = flat.numel()
        nh, step = self._tick(name)
        ns = int(math.ceil(numel * 0.01))
        up = philox_uniform(ns, self.seed, nh, 2 * step, STREAM_SAMPLE, device=flat.device)
        um = philox_uniform(ns, self.seed, nh, 2 * step + 1, STREAM_SAMPLE, device=flat.device)
        pi, pm = _adaq_part(flat, flat > 0, self.compress_ratio, up)
        mi, mm = _adaq_part(flat, flat < 0, self.compress_ratio, um)
        head = torch.stack([pm, mm]).to(torch.float32).view(torch.int32)
        count = torch.tensor([pi.numel()], dtype=torch.int32, device=flat.device)
        wire = torch.cat([head, count, pi.to(torch.int32), mi.to(torch.int32)])
        return [wire], (tensor.size(), numel)

    def decompress(self, tensors, ctx):
        (wire,) = tensors
        shape, numel = ctx
        means = wire[:2].contiguous().view(torch.float32)
        n_plus = int(wire[2])
        out = torch.zeros(numel, dtype=torch.float32, device=wire.device)
        out[wire[3:3 + n_plus].to(torch.int64)] = means[0]
        out[wire[3 + n_plus:].to(torch.int64)] = means[1]
        return out.view(shape)
