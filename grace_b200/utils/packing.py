"""2-bit x 4 -> byte helpers (reference: ``grace_dl/tensorflow/compressor/packing.py:4-30``,
``encode_byte`` / ``decode_byte``; SURVEY §2.3 last row).

The reference keeps these as a stand-alone TF helper that duplicates the inline packing of
INCEPTIONN.  Here both names are thin aliases of the one implementation the INCEPTIONN codec
uses (``grace_b200/compressor/inceptionn.py``): element ``i`` of quarter ``j`` of the (padded)
tag vector lands in bits ``2j..2j+1`` of byte ``i``.  (The native K2-inceptionn kernel keeps
its tags tile-local instead: element ``e`` sits in bits ``2(e%4)`` of byte ``e//4``.)
"""
from __future__ import annotations

import torch

from grace_b200.compressor.inceptionn import pack_tags, unpack_tags


def encode_byte(tags: torch.Tensor) -> torch.Tensor:
    """Pack a 1-D tensor of values in ``0..3`` four to a byte (uint8, ``numel // 4 + 1`` bytes)."""
    return pack_tags(tags.flatten())


def decode_byte(packed: torch.Tensor, numel: int) -> torch.Tensor:
    """Inverse of :func:`encode_byte`: the first ``numel`` 2-bit values as int32."""
    return unpack_tags(packed.flatten(), numel)


__all__ = ["encode_byte", "decode_byte"]
