"""Host-side logic of the fused engine that needs no GPU: triple -> native parameters dispatch,
registry defaults, optimizer wrapper bookkeeping."""
import pytest
import torch

from grace_b200 import compressor as C
from grace_b200 import memory as M
from grace_b200.ops.native import CODEC, MEMORY
from grace_b200.parallel.fused import codec_params, communicator_is_fusable, powersgd_spec


def test_every_builtin_codec_maps_to_a_native_codec_or_is_declared_generic():
    native = {
        C.NoneCompressor(): "none", C.FP16Compressor(): "fp16", C.TopKCompressor(0.01): "topk",
        C.RandomKCompressor(0.01): "randomk", C.ThresholdCompressor(0.1): "threshold", C.DgcCompressor(0.01): "dgc",
        C.QSGDCompressor(64): "qsgd", C.TernGradCompressor(): "terngrad", C.NaturalCompressor(): "natural",
        C.U8bitCompressor(): "u8bit", C.SignSGDCompressor(): "signsgd", C.SignumCompressor(0.9): "signum",
        C.EFSignSGDCompressor(0.1): "efsignsgd", C.OneBitCompressor(): "onebit", C.SketchCompressor(64): "sketch",
        C.AdaqCompressor(0.01): "adaq", C.INCEPTIONNCompressor(): "inceptionn",
    }
    for comp, name in native.items():
        p = codec_params(comp, M.NoneMemory())
        assert p is not None and p.codec == CODEC[name], name
    assert codec_params(C.SketchCompressor(1000), M.NoneMemory()).quantum_num == 1000      # 16-bit bins: native up to 1023
    for comp in (C.SketchCompressor(2000), C.INCEPTIONNCompressor(error_bound=1e300), C.PowerSGDCompressor()):
        assert codec_params(comp, M.NoneMemory()) is None


def test_memory_mapping_and_hyperparameters():
    p = codec_params(C.TopKCompressor(0.01), M.ResidualMemory(beta=0.9, gamma=1.1))
    assert p.memory == MEMORY["residual"] and abs(p.beta - 0.9) < 1e-7 and abs(p.gamma - 1.1) < 1e-7 and p.ratio == 0.01
    p = codec_params(C.EFSignSGDCompressor(0.3), M.EFSignSGDMemory(0.3))
    assert p.memory == MEMORY["residual"] and p.beta == 1.0 and abs(p.gamma - 0.3) < 1e-7 and p.average == 0
    p = codec_params(C.DgcCompressor(0.01), M.DgcMemory(0.9, False, 4))
    assert p.memory == MEMORY["dgc"] and abs(p.momentum - 0.9) < 1e-7
    clipped = codec_params(C.DgcCompressor(0.01), M.DgcMemory(0.9, 0.3, 4))            # the dist helper's default (truthy 0.3)
    assert clipped is not None and clipped.clip == 1                                   # native: scalar all-reduce on the engine
    assert codec_params(C.TopKCompressor(0.01), M.DgcMemory(0.9, False, 4)) is None    # DGC memory needs DGC codec


def test_user_subclasses_take_the_generic_path():
    class MyTopK(C.TopKCompressor):
        pass

    assert codec_params(MyTopK(0.01), M.NoneMemory()) is None


def test_communicator_dispatch():
    from grace_b200.dist.communicator import Allreduce, Allgather, Broadcast
    topk = codec_params(C.TopKCompressor(0.01), M.NoneMemory())
    none = codec_params(C.NoneCompressor(), M.NoneMemory())
    thr = codec_params(C.ThresholdCompressor(0.1), M.NoneMemory())
    assert communicator_is_fusable(Allgather(C.TopKCompressor(0.01), M.NoneMemory(), 2), topk)
    assert not communicator_is_fusable(Allreduce(C.TopKCompressor(0.01), M.NoneMemory(), 2), topk)
    assert communicator_is_fusable(Allreduce(C.NoneCompressor(), M.NoneMemory(), 2), none)
    assert not communicator_is_fusable(Broadcast(C.ThresholdCompressor(0.1), M.NoneMemory(), 2, rank=0), thr)


def test_powersgd_spec():
    from grace_b200.dist import grace_from_params
    g = grace_from_params(dict(compressor="powersgd", memory="powersgd", communicator="allreduce", world_size=2,
                               compress_rank=4))
    assert powersgd_spec(g) == (4, g.compressor.seed, True, False)
    g = grace_from_params(dict(compressor="powersgd", memory="none", communicator="allreduce", world_size=2,
                               compress_rank=8))
    assert powersgd_spec(g) == (8, g.compressor.seed, False, False)                    # ranks up to 16 are native
    g = grace_from_params(dict(compressor="powersgd", memory="powersgd", communicator="allreduce", world_size=2,
                               compress_rank=2, warm_start=True))
    assert powersgd_spec(g) == (2, g.compressor.seed, True, True)                      # warm start: native as well
    g = grace_from_params(dict(compressor="powersgd", memory="none", communicator="allreduce", world_size=2,
                               compress_rank=17))
    assert powersgd_spec(g) is None                                                    # beyond UMMA N = 16: generic path
    g = grace_from_params(dict(compressor="topk", memory="none", communicator="allgather", world_size=2))
    assert powersgd_spec(g) is None


def test_fused_true_raises_on_cpu_parameters():
    import grace_b200.torch as hvd
    m = torch.nn.Linear(4, 2)
    opt = torch.optim.SGD(m.parameters(), lr=0.1)
    grc = hvd.grace_from_params(dict(compressor="topk", memory="residual", communicator="allgather"))
    with pytest.raises(RuntimeError):
        hvd.DistributedOptimizer(opt, grc, named_parameters=m.named_parameters(), fused=True)


def test_optimizer_wrapper_validates_names():
    import grace_b200.torch as hvd
    m = torch.nn.Linear(4, 2)
    grc = hvd.grace_from_params(dict(compressor="none", memory="none", communicator="allreduce"))
    with pytest.raises(ValueError):
        hvd.DistributedOptimizer(torch.optim.SGD(m.parameters(), lr=0.1), grc,
                                 named_parameters=[("w", m.weight), ("w", m.bias)])
    with pytest.raises(ValueError):
        hvd.DistributedOptimizer(torch.optim.SGD(m.parameters(), lr=0.1), grc, named_parameters=[("w", m.weight)])
