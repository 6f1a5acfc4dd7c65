"""The five memories of the reference, restated once for both API flavours.

Reference: ``grace_dl/dist/memory/none.py:4-11``, ``residual.py:4-20``, ``dgc.py:7-39``,
``efsignsgd.py:4-19``, ``powersgd.py:6-37``.  In the fused path this state lives in flat
per-bucket device buffers (``grace_b200/parallel/fused.py``); these classes are the oracle
and serve user-defined combinations.
"""
from __future__ import annotations

import torch
from torch import distributed as dist

from grace_b200.core import Memory
from grace_b200.compressor.lowrank import RANK_KEY


class NoneMemory(Memory):
    def compensate(self, tensor, name):
        return tensor


class ResidualMemory(Memory):
    """``x <- beta * e + gamma * g``;  ``e <- x - decompress(compress(x))``."""

    _state_fields = ("residuals",)

    def __init__(self, beta=1.0, gamma=1.0):
        self.residuals = {}
        self.beta = beta
        self.gamma = gamma

    def compensate(self, tensor, name):
        if name in self.residuals:
            tensor = self.beta * self.residuals[name] + self.gamma * tensor
        return tensor

    def update(self, tensor, name, compressor, tensor_compressed, ctx):
        self.residuals[name] = tensor - compressor.decompress(tensor_compressed, ctx)


class EFSignSGDMemory(Memory):
    """``x <- e + lr * g`` (reference ``memory/efsignsgd.py:9-13``)."""

    _state_fields = ("residuals",)

    def __init__(self, lr):
        self.residuals = {}
        self.learning_rate = lr

    def compensate(self, tensor, name):
        if name in self.residuals:
            tensor = self.residuals[name] + self.learning_rate * tensor
        return tensor

    def update(self, tensor, name, compressor, tensor_compressed, ctx):
        self.residuals[name] = tensor - compressor.decompress(tensor_compressed, ctx)


class DgcMemory(Memory):
    """Momentum correction + local accumulation, masked by what DGC sent.

    ``u <- m u + g``, ``v <- v + u``; after selection both are zeroed where the
    compressor's mask is set.  Optional clipping to ``sqrt(allreduce(sum g^2) / W)`` --
    the formula the reference intends (its dist flavour crashes there)."""

    _state_fields = ("residuals", "gradients")

    def __init__(self, momentum, gradient_clipping=False, world_size=1, allreduce_sum=None):
        self.momentum = momentum
        self.gradient_clipping = gradient_clipping
        self.world_size = world_size
        self.residuals = {}
        self.gradients = {}
        self._allreduce_sum = allreduce_sum

    def _sum_over_ranks(self, t):
        if self._allreduce_sum is not None:
            return self._allreduce_sum(t)
        if self.world_size > 1 and dist.is_available() and dist.is_initialized():
            dist.all_reduce(t)
        return t

    def compensate(self, tensor, name):
        if self.gradient_clipping:
            sq = self._sum_over_ranks(torch.sum(tensor * tensor))
            clip = torch.sqrt(sq / self.world_size)
            tensor = torch.maximum(torch.minimum(tensor, clip), -clip)
        if name in self.residuals:
            self.residuals[name] = self.momentum * self.residuals[name] + tensor
        else:
            self.residuals[name] = tensor.clone()
        if name in self.gradients:
            self.gradients[name] = self.gradients[name] + self.residuals[name]
        else:
            self.gradients[name] = tensor.clone()
        return self.gradients[name]

    def update(self, tensor, name, compressor, tensor_compressed, ctx):
        shape, mask, _ = ctx
        keep = ~mask.view(shape)
        self.residuals[name] = self.residuals[name] * keep
        self.gradients[name] = self.gradients[name] * keep


class PowerSGDMemory(Memory):
    """Residual for matrices; shares ``q_memory`` with ``PowerSGDCompressor``.

    Like the reference (``memory/powersgd.py:20-28``) the default drops the stored Q every
    step so the compressor draws a fresh one; ``warm_start=True`` keeps it."""

    _state_fields = ("residuals",)

    def __init__(self, q_memory, compress_rank=1, warm_start=False):
        self.q_memory = q_memory
        self.compress_rank = compress_rank
        self.warm_start = warm_start
        self.residuals = {}
        self.q_memory[RANK_KEY] = compress_rank

    def compensate(self, tensor, name):
        if tensor.dim() == 1:
            return tensor
        if name in self.residuals:
            tensor = tensor + self.residuals[name]
        if not self.warm_start:
            self.q_memory.pop(name, None)
        return tensor

    def update(self, tensor, name, compressor, tensor_compressed, ctx):
        if ctx is None:
            return
        self.residuals[name] = tensor - compressor.decompress(tensor_compressed, ctx)
