"""Observability: CUDA-event timers, NVTX ranges and per-step compression counters.

The reference has only wall-clock prints around ``torch.cuda.synchronize()``
(``examples/torch/pytorch_mnist.py:148-155``, ``grace_dl/torch/compressor/qsgd.py:14-15``) and no
bytes-on-wire or residual counters.  Everything here is opt-in and off the hot path:
nothing synchronises the device unless ``summary()`` / ``elapsed_ms()`` is called.
"""
from __future__ import annotations

import contextlib
import json
import time
from collections import defaultdict

import torch


@contextlib.contextmanager
def nvtx(name):
    """NVTX range visible in Nsight tools (no-op without CUDA)."""
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class CudaTimer:
    """Device-side timing with CUDA events on the current stream; host time on CPU."""

    def __init__(self):
        self.records = defaultdict(list)
        self._open = {}

    def start(self, name):
        if torch.cuda.is_available():
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self._open[name] = ev
        else:
            self._open[name] = time.perf_counter()

    def stop(self, name):
        begin = self._open.pop(name)
        if torch.cuda.is_available():
            end = torch.cuda.Event(enable_timing=True)
            end.record()
            self.records[name].append((begin, end))
        else:
            self.records[name].append((time.perf_counter() - begin) * 1e3)

    @contextlib.contextmanager
    def section(self, name):
        self.start(name)
        with nvtx(name):
            yield
        self.stop(name)

    def elapsed_ms(self):
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        out = {}
        for name, recs in self.records.items():
            vals = [r if isinstance(r, float) else r[0].elapsed_time(r[1]) for r in recs]
            out[name] = {"n": len(vals), "mean_ms": sum(vals) / len(vals), "max_ms": max(vals)}
        return out


class CompressionStats:
    """Counters for a ``DistributedOptimizer`` running the fused engine."""

    def __init__(self, optimizer):
        self.fused = getattr(optimizer, "_fused", None)
        self.steps = 0
        self._launch0 = self.fused.launches() if self.fused else 0
        self.graph_launches_per_step = None     # set to GraphedStep.native_launches_per_step: replays bypass the host counter

    def step(self):
        self.steps += 1

    def summary(self, step_ms=None):
        if self.fused is None:
            return {"fused": False}
        dense = sum(p.numel() * 4 for b in self.fused.buckets for p in b.tensors)
        wire = self.fused.wire_bytes()
        world = self.fused.engine.world
        out = {"fused": True, "buckets": len(self.fused.buckets), "world": world, "dense_gradient_bytes": dense,
               "wire_bytes_per_peer_per_step": wire, "compression_ratio": dense / max(wire, 1),
               "egress_bytes_per_step": wire * max(world - 1, 0),
               "launches_per_step": self.graph_launches_per_step if self.graph_launches_per_step is not None
               else (self.fused.launches() - self._launch0) / max(self.steps, 1)}
        if step_ms:
            out["egress_GBps"] = out["egress_bytes_per_step"] / step_ms / 1e6
        return out

    def residual_norms(self):
        """L2 norm of every bucket's error-feedback buffer (synchronises the device)."""
        norms = []
        for b in self.fused.buckets if self.fused else []:
            for pl in b.plans:
                s = pl.get_state(0)
                norms.append(float(s.norm()) if s is not None else 0.0)
        return norms


class TableLogger:
    """Fixed-width table on rank 0 + optional TSV/JSON lines file (reference
    ``examples/dist/CIFAR10-dawndist/core.py:33-39``, ``dawn.py:72-81``)."""

    def __init__(self, path=None, rank=0):
        self.keys, self.path, self.rank = None, path, rank

    def append(self, row):
        if self.rank != 0:
            return
        if self.keys is None:
            self.keys = list(row)
            print(*("%12s" % k for k in self.keys))
        print(*("%12.4f" % row[k] if isinstance(row[k], float) else "%12s" % row[k] for k in self.keys))
        if self.path:
            with open(self.path, "a") as f:
                f.write(json.dumps(row) + "\n")


class TSVLogger:
    """DAWNBench submission log: ``epoch<TAB>hours<TAB>top1Accuracy`` per epoch (reference ``dawn.py:72-81``; written to
    ``logs.tsv`` at the end of training, ``dawn.py:132-133``)."""

    def __init__(self):
        self.log = ["epoch\thours\ttop1Accuracy"]

    def append(self, row):
        self.log.append("%d\t%.8f\t%.2f" % (row["epoch"], row["total_time"] / 3600.0, row["test_acc"] * 100.0))

    def __str__(self):
        return "\n".join(self.log)

    def write(self, path):
        with open(path, "w") as f:
            f.write(str(self))


class Timer:
    """Wall-clock timer that synchronises the device before every reading (reference ``core.py:14-27``):
    ``timer()`` returns the time since the previous call and adds it to ``total_time`` unless told otherwise --
    DAWNBench excludes evaluation (``test_time_in_total=False``, ``dawn.py:129``)."""

    def __init__(self, synch=None):
        import time
        self._time = time.time
        self.synch = synch or (lambda: None)
        self.synch()
        self.times = [self._time()]
        self.total_time = 0.0

    def __call__(self, include_in_total=True):
        self.synch()
        self.times.append(self._time())
        delta = self.times[-1] - self.times[-2]
        if include_in_total:
            self.total_time += delta
        return delta
