"""``topk`` module of the reference (``grace_dl/dist/compressor/topk.py``): the class plus its two
module-level helpers, which user code imports directly."""
import torch

from grace_b200.compressor import TopKCompressor  # noqa: F401
from grace_b200.compressor._common import scatter_dense, topk_count


def sparsify(tensor, compress_ratio):
    """``(values, indices)`` of the ``max(1, int(numel * ratio))`` largest-magnitude entries (topk.py:6-11)."""
    flat = tensor.flatten()
    _, indices = torch.topk(flat.abs(), topk_count(flat.numel(), compress_ratio), sorted=False)
    return flat[indices], indices


def desparsify(tensors, numel):
    """Dense 1-D tensor of ``numel`` entries with ``values`` scattered at ``indices`` (topk.py:14-18)."""
    values, indices = tensors
    return scatter_dense(values, indices, numel)
