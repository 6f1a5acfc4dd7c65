"""``powersgd`` module of the reference (``grace_dl/dist/compressor/powersgd.py``): the class plus the
module-level Gram-Schmidt helper (here with an epsilon guard and a second sweep)."""
from grace_b200.compressor import PowerSGDCompressor  # noqa: F401
from grace_b200.compressor.lowrank import orthogonalize  # noqa: F401
