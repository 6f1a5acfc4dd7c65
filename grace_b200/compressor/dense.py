"""Identity and half-precision codecs (reference ``grace_dl/dist/compressor/none.py:4-12``,
``fp16.py:6-22``).  Linear payloads: valid with every communicator."""
from __future__ import annotations

import torch

from grace_b200.core import Compressor


class NoneCompressor(Compressor):
    """Pass-through.  Extra positional arguments are tolerated because the reference
    examples pass stray ones (``examples/torch/pytorch_mnist.py:122``)."""

    def __init__(self, *_ignored, **kw):
        super().__init__(**{k: v for k, v in kw.items() if k in ("average", "tensors_size_are_same")})

    def compress(self, tensor, name):
        return [tensor], None

    def decompress(self, tensors, ctx):
        (tensor,) = tensors
        return tensor


class FP16Compressor(Compressor):
    """Floating tensors travel as float16 and are cast back on receipt."""

    def __init__(self, wire_dtype=torch.float16):
        super().__init__()
        self.wire_dtype = wire_dtype

    def compress(self, tensor, name):
        orig = tensor.dtype
        if orig.is_floating_point:
            tensor = tensor.to(self.wire_dtype)
        return [tensor], orig

    def decompress(self, tensors, ctx):
        (tensor,) = tensors
        return tensor.to(ctx) if ctx.is_floating_point else tensor
