"""Name -> constructor tables behind both ``grace_from_params`` flavours.

The reference resolves names with an if/elif ladder that must be edited to add a codec
(``grace_dl/dist/helper.py:5-55``); here it is a registry that users can extend with
``register_compressor`` / ``register_memory``.
"""
from __future__ import annotations

from grace_b200 import compressor as C
from grace_b200.compressor._common import StochasticCompressor
from grace_b200 import memory as M

# name -> (class, [(ctor kwarg, params key, dist default, torch default)])
COMPRESSORS = {
    "none": (C.NoneCompressor, []),
    "fp16": (C.FP16Compressor, []),
    "topk": (C.TopKCompressor, [("compress_ratio", "compress_ratio", 0.3, 0.01)]),
    "randomk": (C.RandomKCompressor, [("compress_ratio", "compress_ratio", 0.3, 0.01)]),
    "threshold": (C.ThresholdCompressor, [("threshold", "threshold", 256, 0.01)]),
    "dgc": (C.DgcCompressor, [("compress_ratio", "compress_ratio", 0.3, 0.01)]),
    "adaq": (C.AdaqCompressor, [("compress_ratio", "compress_ratio", 0.01, 0.01)]),
    "qsgd": (C.QSGDCompressor, [("quantum_num", "quantum_num", 256, 64)]),
    "terngrad": (C.TernGradCompressor, []),
    "natural": (C.NaturalCompressor, []),
    "u8bit": (C.U8bitCompressor, []),
    "8bit": (C.U8bitCompressor, []),
    "sketch": (C.SketchCompressor, [("quantiles", "quantiles", 64, 64)]),
    "inceptionn": (C.INCEPTIONNCompressor, [("error_bound", "error_bound", 2e-10, 2e-10)]),
    "signsgd": (C.SignSGDCompressor, []),
    "signum": (C.SignumCompressor, [("momentum", "momentum", 0.9, 0.9)]),
    "efsignsgd": (C.EFSignSGDCompressor, [("lr", "lr", 0.3, 0.1)]),
    "onebit": (C.OneBitCompressor, []),
    "powersgd": (C.PowerSGDCompressor, [("rank", "compress_rank", 1, 1)]),
}

MEMORIES = {
    "none": (M.NoneMemory, []),
    "residual": (M.ResidualMemory, [("beta", "beta", 1.0, 1.0), ("gamma", "gamma", 1.0, 1.0)]),
    "efsignsgd": (M.EFSignSGDMemory, [("lr", "lr", 0.3, 0.1)]),
    "dgc": (M.DgcMemory, [("momentum", "momentum", 0.3, 0.9),
                          ("gradient_clipping", "gradient_clipping", 0.3, False)]),
    "powersgd": (M.PowerSGDMemory, [("compress_rank", "compress_rank", 1, 1)]),
}


def register_compressor(name, cls, arguments=()):
    COMPRESSORS[name] = (cls, list(arguments))


def register_memory(name, cls, arguments=()):
    MEMORIES[name] = (cls, list(arguments))


def _kwargs(spec, params, flavour):
    col = 2 if flavour == "dist" else 3
    return {row[0]: params.get(row[1], row[col]) for row in spec}


def build_compressor(params, flavour, world_size):
    name = params.get("compressor", "none")
    if name not in COMPRESSORS:
        raise NotImplementedError(name)
    cls, spec = COMPRESSORS[name]
    kw = _kwargs(spec, params, flavour)
    if cls is C.PowerSGDCompressor:
        kw["world_size"] = world_size                 # form the mean
        kw["warm_start"] = bool(params.get("warm_start", False))
    if "seed" in params and issubclass(cls, StochasticCompressor):
        kw["seed"] = params["seed"]
    return cls(**kw)


def build_memory(params, flavour, world_size, compressor):
    name = params.get("memory", "none")
    if name not in MEMORIES:
        raise NotImplementedError(name)
    cls, spec = MEMORIES[name]
    kw = _kwargs(spec, params, flavour)
    if cls is M.DgcMemory:
        kw["world_size"] = world_size
    if cls is M.PowerSGDMemory:
        if not hasattr(compressor, "q_memory"):
            raise ValueError("memory 'powersgd' needs compressor 'powersgd'")
        return cls(compressor.q_memory, warm_start=bool(params.get("warm_start", False)), **kw)
    return cls(**kw)
