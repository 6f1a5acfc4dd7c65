"""Transport helpers for the generic (collective-library) path."""
import torch

_NATIVE = {torch.float32, torch.float64, torch.float16, torch.int8, torch.uint8, torch.int32, torch.int64}


def as_wire(t):
    """Return ``(buffer, restore)``: dtypes some backends cannot move (int16, bool, bf16 on
    Gloo) travel as raw bytes and are re-viewed on arrival."""
    if t.dtype in _NATIVE:
        return t, (lambda x: x)
    dtype = t.dtype
    if dtype == torch.bool:
        return t.to(torch.uint8), (lambda x: x.to(torch.bool))
    return t.contiguous().view(-1).view(torch.uint8), (lambda x: x.view(dtype))
