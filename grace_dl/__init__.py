"""Import-path compatibility with the reference package (``grace_dl/__init__.py:1``):
``grace_dl.dist`` and ``grace_dl.torch`` resolve to ``grace_b200`` so existing user code
(``from grace_dl.dist.compressor.topk import TopKCompressor`` ...) keeps working."""
__version__ = "1.0"
