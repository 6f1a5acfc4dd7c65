"""Asynchronous communicators over the handle API
(reference ``grace_dl/torch/communicator/allreduce.py:5-15``, ``allgather.py:7-52``,
``broadcast.py:5-27``)."""
from __future__ import annotations

import torch

from grace_b200.core import AsyncCommunicator
from grace_b200.torch import mpi_ops as ops


class _NativeAsync:
    """Mixin: ``send_step`` / ``receive_step`` of built-in triples on CUDA fp32 tensors run on
    single-tensor native plans (compress + peer stores on a side stream at send time, decode +
    aggregate at receive time); everything else takes the generic handle-based path."""

    fused = "auto"
    _stepper = None

    def send_step(self, tensor, name):
        if self.fused and tensor.is_cuda:
            if self._stepper is None:
                from grace_b200.parallel.fused import FusedStepper
                self._stepper = FusedStepper(self)
            ticket = self._stepper.send(tensor, name)
            if ticket is not None:
                return ticket, None
        return super().send_step(tensor, name)

    def receive_step(self, handles, ctx):
        if isinstance(handles, tuple) and handles and handles[0] == "native":
            return self._stepper.receive(handles)
        return super().receive_step(handles, ctx)


def _finish(compressor, decompressed, world_size):
    out = compressor.aggregate(decompressed)
    return out / world_size if compressor.average else out


class Allreduce(_NativeAsync, AsyncCommunicator):
    def async_send(self, tensors, name):
        return [ops.allreduce_async_(t.contiguous(), self.compressor.average, "%s%d" % (name, i))
                for i, t in enumerate(tensors)]

    def wait_receive(self, handles, ctx):
        return self.compressor.decompress([ops.synchronize(h) for h in handles], ctx)


class Allgather(_NativeAsync, AsyncCommunicator):
    def __init__(self, compressor, memory, world_size=None):
        super().__init__(compressor, memory)
        self.world_size = ops.size() if world_size is None else world_size

    def async_send(self, tensors, name):
        sent = []
        for t in tensors:
            rows = t.shape[0] if t.dim() > 0 else 1
            if self.compressor.tensors_size_are_same:
                sizes = [rows] * self.world_size
            else:
                sizes = [int(v) for v in ops.allgather(torch.tensor([rows], dtype=torch.int64, device=t.device))]
            sent.append((ops.allgather_async(t), sizes, t.dim() == 0))
        return sent

    def wait_receive(self, result, ctx):
        columns = []
        for handle, sizes, scalar in result:
            parts = ops.synchronize(handle).split(sizes)
            columns.append([p.reshape(()) for p in parts] if scalar else list(parts))
        if not columns:                      # empty payload (e.g. PowerSGD matrices)
            per_rank = [[] for _ in range(self.world_size)]
        else:
            per_rank = [list(col) for col in zip(*columns)]
        decompressed = [self.compressor.decompress(p, ctx) for p in per_rank]
        return _finish(self.compressor, decompressed, self.world_size)


class Broadcast(_NativeAsync, AsyncCommunicator):
    def __init__(self, compressor, memory, world_size=None):
        super().__init__(compressor, memory)
        self.world_size = ops.size() if world_size is None else world_size

    def async_send(self, tensors, name):
        if not self.compressor.tensors_size_are_same:
            raise NotImplementedError("Broadcast cannot carry variable-size payloads; use Allgather")
        return [[ops.broadcast_async(t, root, "%s%d_%d" % (name, root, i)) for i, t in enumerate(tensors)]
                for root in range(self.world_size)]

    def wait_receive(self, handles, ctx):
        decompressed = [self.compressor.decompress([ops.synchronize(h) for h in per_root], ctx)
                        for per_root in handles]
        return _finish(self.compressor, decompressed, self.world_size)
