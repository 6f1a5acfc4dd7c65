"""API parity with the reference, file by file: every public class / function defined in the reference's
``grace_dl/{dist,torch}`` modules must be importable under the same module path and name from this repo's
``grace_dl`` shim, and the Horovod-style module must expose the reference's ``mpi_ops`` surface."""
import ast
import importlib
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = ["/root/reference/grace_dl", os.path.join(ROOT, "baseline", "_ref", "grace_dl")]
REF = next((p for p in CANDIDATES if os.path.isdir(os.path.join(p, "dist"))), None)

pytestmark = pytest.mark.skipif(REF is None, reason="reference sources not available")


def _public_names(path):
    tree = ast.parse(open(path).read())
    return [n.name for n in tree.body if isinstance(n, (ast.ClassDef, ast.FunctionDef)) and not n.name.startswith("_")]


def _modules():
    out = []
    for flavour in ("dist", "torch"):
        base = os.path.join(REF, flavour)
        for dirpath, _, files in os.walk(base):
            for f in sorted(files):
                if f.endswith(".py"):
                    rel = os.path.relpath(os.path.join(dirpath, f), REF)[:-3].replace(os.sep, ".")
                    if rel.endswith(".__init__"):
                        rel = rel[:-9]
                    out.append(("grace_dl." + rel, os.path.join(dirpath, f)))
    return out


@pytest.mark.parametrize("module,path", _modules() if REF else [], ids=[m for m, _ in (_modules() if REF else [])])
def test_reference_module_names_are_importable(module, path):
    names = _public_names(path)
    mod = importlib.import_module(module)
    import grace_dl
    assert os.path.realpath(grace_dl.__file__).startswith(os.path.realpath(ROOT) + os.sep + "grace_dl"), grace_dl.__file__
    missing = [n for n in names if not hasattr(mod, n)]
    assert not missing, "%s lacks %s" % (module, missing)


def test_horovod_style_surface():
    """Names the reference's patched ``horovod.torch`` exports (patch_files/horovod/torch/__init__.py:23-30,
    mpi_ops.py) and its examples call."""
    import grace_b200.torch as hvd
    for name in ["init", "shutdown", "size", "local_size", "rank", "local_rank", "mpi_threads_supported",
                 "allreduce", "allreduce_async", "allreduce_", "allreduce_async_", "allgather", "allgather_async",
                 "broadcast", "broadcast_async", "broadcast_", "broadcast_async_", "poll", "synchronize",
                 "DistributedOptimizer", "broadcast_parameters", "broadcast_optimizer_state", "grace_from_params",
                 "Memory", "Compressor", "Communicator"]:
        assert hasattr(hvd, name), name


def test_constructor_signatures_match_the_reference():
    """Positional constructor arguments used by the reference's examples keep working."""
    from grace_dl.dist.communicator.allgather import Allgather
    from grace_dl.dist.communicator.allreduce import Allreduce
    from grace_dl.dist.compressor.topk import TopKCompressor
    from grace_dl.dist.compressor.qsgd import QSGDCompressor
    from grace_dl.dist.compressor.powersgd import PowerSGDCompressor
    from grace_dl.dist.memory.residual import ResidualMemory
    from grace_dl.dist.memory.dgc import DgcMemory
    from grace_dl.dist.memory.powersgd import PowerSGDMemory
    g = Allgather(TopKCompressor(0.3), ResidualMemory(), 2)
    assert g.world_size == 2 and g.compressor.compress_ratio == 0.3
    Allreduce(QSGDCompressor(64), ResidualMemory(1.0, 1.0), 2)
    DgcMemory(0.9, False, 2)
    c = PowerSGDCompressor()
    PowerSGDMemory(c.q_memory, 1)


def test_dawn_augmentation_matches_the_reference_transforms():
    """Crop(32, 32) + FlipLR + Cutout(8, 8) of examples/dist/CIFAR10-dawndist/core.py:69-110, batched on the device:
    every output is a 32x32 window of its padded source (possibly mirrored) with one 8x8 hole of zeros."""
    import torch
    from grace_b200.utils.augment import DawnAugment, normalise, pad_reflect
    g = torch.Generator().manual_seed(0)
    raw = torch.randint(1, 255, (16, 32, 32, 3), dtype=torch.uint8, generator=g)
    x = normalise(raw)
    assert x.shape == (16, 3, 32, 32) and abs(float(x.mean())) < 1.0
    padded = pad_reflect(x, 4)
    assert padded.shape == (16, 3, 40, 40) and torch.equal(padded[:, :, 4:36, 4:36], x)
    out = DawnAugment(generator=g)(padded)
    assert out.shape == (16, 3, 32, 32)
    for i in range(16):
        hole = (out[i] == 0).all(0)
        assert int(hole.sum()) == 64                               # one 8x8 cut-out
        found = False
        for flip in (False, True):
            src = padded[i].flip(2) if flip else padded[i]
            windows = src.unfold(1, 32, 1).unfold(2, 32, 1)        # [3, 9, 9, 32, 32]
            match = ((windows == out[i].view(3, 1, 1, 32, 32)) | hole.view(1, 1, 1, 32, 32)).all(-1).all(-1).all(0)
            found = found or bool(match.any())
        assert found


def test_tsv_logger_and_timer():
    from grace_b200.utils.metrics import Timer, TSVLogger
    t = Timer()
    a = t()
    b = t(include_in_total=False)
    assert a >= 0 and b >= 0 and abs(t.total_time - a) < 1e-9
    log = TSVLogger()
    log.append({"epoch": 3, "total_time": 7200.0, "test_acc": 0.9412})
    assert str(log).splitlines() == ["epoch\thours\ttop1Accuracy", "3\t2.00000000\t94.12"]


def test_capture_hint_lifts_the_mask_guard_only_for_maskless_calls():
    """grace_b200.utils.capture: inside the context transformers' `is_tracing` guard answers False for a call without a
    padding mask (so no all-ones mask is materialised while a stream captures) and is untouched for calls with one;
    the original function is restored on exit."""
    import torch
    mu = pytest.importorskip("transformers.masking_utils")
    from grace_b200.utils.capture import unmasked_attention_when_capturing
    orig = mu.is_tracing
    sentinel = torch.ones(2, 4)
    with unmasked_attention_when_capturing():
        assert mu.is_tracing is not orig
        assert mu.is_tracing(None) is False and mu.is_tracing() is False
        assert mu.is_tracing(sentinel) == orig(sentinel)
    assert mu.is_tracing is orig
