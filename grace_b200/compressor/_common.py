"""Helpers shared by the torch restatements of the codecs."""
from __future__ import annotations

import torch

from grace_b200.core import Compressor
from grace_b200.utils.philox import fnv1a

DEFAULT_SEED = 0x9E3779B9 & 0x7FFFFFFF


class StochasticCompressor(Compressor):
    """Base for codecs that draw random numbers.  Draws come from the Philox stream
    keyed by ``(seed, fnv1a(name))`` with a per-name call counter, so they are
    reproducible, rank-invariant when the seed is shared, and identical to what the
    fused CUDA kernels consume (``grace_b200/utils/philox.py``)."""

    _state_fields = ("steps",)

    def __init__(self, average=True, tensors_size_are_same=True, seed=DEFAULT_SEED):
        super().__init__(average=average, tensors_size_are_same=tensors_size_are_same)
        self.seed = int(seed) & 0xFFFFFFFF
        self.steps = {}

    def _tick(self, name):
        """Return ``(name_hash, step)`` for this call and advance the counter."""
        step = self.steps.get(name, 0)
        self.steps[name] = step + 1
        return fnv1a(name), step


def scatter_dense(values, indices, numel):
    out = torch.zeros(numel, dtype=values.dtype, device=values.device)
    out.scatter_(0, indices.to(torch.int64), values)
    return out


def topk_count(numel, ratio):
    """``k = max(1, int(numel * ratio))`` -- reference ``compressor/topk.py:8``."""
    return max(1, int(numel * ratio))
