"""Compressor zoo.  One implementation serves ``grace_b200.dist`` and
``grace_b200.torch``."""
from grace_b200.compressor.dense import NoneCompressor, FP16Compressor  # noqa: F401
from grace_b200.compressor.sparse import (  # noqa: F401
    TopKCompressor, RandomKCompressor, ThresholdCompressor, DgcCompressor, AdaqCompressor)
from grace_b200.compressor.quant import (  # noqa: F401
    QSGDCompressor, TernGradCompressor, NaturalCompressor, U8bitCompressor, SketchCompressor)
from grace_b200.compressor.sign import (  # noqa: F401
    SignSGDCompressor, SignumCompressor, EFSignSGDCompressor, OneBitCompressor)
from grace_b200.compressor.lowrank import PowerSGDCompressor  # noqa: F401
from grace_b200.compressor.inceptionn import INCEPTIONNCompressor  # noqa: F401
