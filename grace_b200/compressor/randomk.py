"""``randomk`` module of the reference (``grace_dl/dist/compressor/randomk.py``): the class plus its
module-level helper."""
import torch

from grace_b200.compressor import RandomKCompressor  # noqa: F401
from grace_b200.compressor._common import topk_count


def sparsify(tensor, compress_ratio):
    """``(values, indices)`` at ``k`` uniformly random distinct positions drawn from torch's global generator, like
    the reference helper (randomk.py:6-12).  ``RandomKCompressor`` itself uses a private keyed permutation and does
    not touch the global generator."""
    flat = tensor.flatten()
    k = topk_count(flat.numel(), compress_ratio)
    indices = torch.randperm(flat.numel(), device=flat.device)[:k]
    return flat[indices], indices
