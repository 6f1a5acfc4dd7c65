"""One-bit codecs: SignSGD, Signum, EFSignSGD, OneBit.

Reference files: ``grace_dl/dist/compressor/signsgd.py:6-30``, ``signum.py:6-37``,
``efsignsgd.py:6-33`` and ``grace_dl/torch/compressor/onebit.py:6-32`` (the torch
flavour; the dist flavour's ``~mask0`` on uint8 is a known defect).
``average=False`` as in the torch/dist flavours.

The Python-visible payload is one uint8 per element like the reference; the fused path
(``csrc/k2_quant.cu``) ballots 32 signs into one word and votes with popcounts.
"""
from __future__ import annotations

import torch

from grace_b200.core import Compressor


def _majority(tensors):
    """Element-wise majority vote of +-1 tensors, tie -> +1 (``signsgd.py:25-30``)."""
    total = sum(tensors)
    return (total >= 0).to(torch.float32) * 2.0 - 1.0


def _unsign(bits, shape):
    return (bits.to(torch.float32) * 2 - 1).view(shape)


class SignSGDCompressor(Compressor):
    def __init__(self):
        super().__init__(average=False)

    def compress(self, tensor, name):
        return [(tensor.flatten() >= 0).to(torch.uint8)], tensor.size()

    def decompress(self, tensors, ctx):
        (bits,) = tensors
        return _unsign(bits, ctx)

    def aggregate(self, tensors):
        return _majority(tensors)


class SignumCompressor(Compressor):
    """SignSGD on a per-name exponential moving average of the gradient."""

    _state_fields = ("momentums",)

    def __init__(self, momentum):
        super().__init__(average=False)
        self.momentum = momentum
        self.momentums = {}

    def compress(self, tensor, name):
        flat = tensor.flatten()
        if name in self.momentums:
            flat = (1.0 - self.momentum) * flat + self.momentum * self.momentums[name]
        else:
            # the reference keeps this view (signum.py:19-23): its own training loop then overwrites the gradient
            # with the decompressed signs and zeroes it, which silently replaces the first momentum by +-1 / 0
            flat = flat.clone()
        self.momentums[name] = flat
        return [(flat >= 0).to(torch.uint8)], tensor.size()

    def decompress(self, tensors, ctx):
        (bits,) = tensors
        return _unsign(bits, ctx)

    def aggregate(self, tensors):
        return _majority(tensors)


class EFSignSGDCompressor(Compressor):
    """Scaled sign: ``mean|x| * sign(x)``; the aggregate is divided by ``lr`` to undo
    the scaling applied by ``EFSignSGDMemory``."""

    def __init__(self, lr):
        super().__init__(average=False)
        self.learning_rate = lr

    def compress(self, tensor, name):
        flat = tensor.flatten()
        return [flat.abs().mean(), (flat >= 0).to(torch.uint8)], tensor.size()

    def decompress(self, tensors, ctx):
        mean, bits = tensors
        return (mean * (bits.to(torch.float32) * 2 - 1)).view(ctx)

    def aggregate(self, tensors):
        return sum(tensors) / self.learning_rate


class OneBitCompressor(Compressor):
    """1-bit SGD: negative entries are replaced by their mean, non-negative by theirs."""

    def compress(self, tensor, name):
        flat = tensor.flatten()
        neg = flat < 0
        n_neg = neg.sum().to(flat.dtype)
        n_pos = flat.numel() - n_neg
        s_neg = torch.where(neg, flat, torch.zeros_like(flat)).sum()
        s_pos = torch.where(neg, torch.zeros_like(flat), flat).sum()
        mean_neg = torch.where(n_neg > 0, s_neg / n_neg.clamp(min=1), s_neg)
        mean_pos = torch.where(n_pos > 0, s_pos / n_pos.clamp(min=1), s_pos)
        return [neg.to(torch.uint8), mean_neg, mean_pos], tensor.size()

    def decompress(self, tensors, ctx):
        neg, mean_neg, mean_pos = tensors
        neg = neg.bool()
        return torch.where(neg, mean_neg, mean_pos).view(ctx)
