"""Helpers for models that go into a whole-step CUDA graph (``grace_b200.torch.GraphedStep``).

A captured step replays exactly the kernels the eager code chose WHILE the stream was capturing, and some model
libraries choose differently then.  Measured case (``benchmarks/graph_vs_eager.py``, BERT-large, batch 32 x 128):
Hugging Face transformers will not look at the contents of a padding mask while a stream is capturing
(``transformers/masking_utils.py``: ``_ignore_bidirectional_mask_sdpa`` is guarded by ``is_tracing``), so even with
``attention_mask=None`` it materialises an all-ones 4-D fp32 mask; with a mask, ``scaled_dot_product_attention`` leaves
the cuDNN flash kernel for the math path, which up-casts bf16 to fp32: 3 757 kernels / 42.5 ms of kernel time per
step instead of 2 956 / 31.2 ms eager -- the graph replays slower (41.1 ms) than the eager loop (35.8 ms) although it
removes all launch gaps.  With no mask there is nothing data-dependent to look at, so the guard can be lifted.
"""
from __future__ import annotations

import contextlib


@contextlib.contextmanager
def unmasked_attention_when_capturing():
    """Inside this context, transformers skips building a bidirectional / causal SDPA mask when the caller passed
    ``attention_mask=None`` even while a CUDA stream is capturing (so SDPA keeps its fused kernel in the graph).
    Calls that do pass a mask are untouched.  No-op when transformers is not importable."""
    try:
        import transformers.masking_utils as mu
    except Exception:                       # pragma: no cover - transformers not installed
        yield
        return
    orig = mu.is_tracing

    def is_tracing(tensor=None):
        return tensor is not None and orig(tensor)

    mu.is_tracing = is_tracing
    try:
        yield
    finally:
        mu.is_tracing = orig
