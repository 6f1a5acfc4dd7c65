"""Generic (any codec, any device, any backend) communicators over ``torch.distributed``.

Reference: ``grace_dl/dist/communicator/allreduce.py:6-13``, ``allgather.py:7-45``,
``broadcast.py:7-33``.  This is the plumbing path -- CPU/Gloo, user-defined codecs and the
oracle for the fused engine.  Built-in (memory, codec, communicator) triples on CUDA tensors
are normally executed by ``grace_b200.parallel`` as one send kernel + one receive kernel
with no collective-library call at all.

Changes against the reference: variable-size gathers exchange the size vector on the
tensor's own device; ``Broadcast`` takes its rank from the process
group when not given (item 6); empty payload lists (PowerSGD) are legal for ``Allreduce``.
"""
from __future__ import annotations

import torch
from torch import distributed as dist

from grace_b200.core import Communicator
from grace_b200.utils.wire import as_wire


def _live(world_size):
    return world_size > 1 and dist.is_available() and dist.is_initialized()


class _NativeStep:
    """Mixin: CUDA tensors of built-in triples take the native path (one single-segment plan
    per tensor name, ``grace_b200/parallel/fused.py``); everything else runs the generic
    pipeline of the base class.  ``fused=False`` (or ``GRACE_B200_FUSED=0``) disables it."""

    fused = "auto"
    _stepper = None

    def step(self, tensor, name):
        if self.fused and tensor.is_cuda:
            if self._stepper is None:
                from grace_b200.parallel.fused import FusedStepper
                self._stepper = FusedStepper(self)
            out = self._stepper.step(tensor, name)
            if out is not None:
                return out
        return super().step(tensor, name)

    def step_all(self, named_grads, inplace=False):
        """Bucketed form of the reference's manual loop (``examples/dist/CIFAR10-dawndist/core.py:202-208``:
        ``for name, p in named_parameters(): grad.copy_(grc.step(grad, name))``): exchange every ``(name, grad)`` pair
        of the list in one call.  On CUDA, built-in triples run as O(#buckets) native launches (the per-tensor ``k``,
        norms, residuals are preserved by the segment tables) instead of O(#tensors x #ops); anything else falls back
        to ``step`` per tensor.  Returns the aggregated tensors in order; with ``inplace=True`` they are also copied
        into the given gradient tensors (one multi-tensor copy).  On the native path the returned tensors are views of
        persistent bucket memory and are overwritten by the next ``step_all``."""
        named_grads = list(named_grads)
        outs = None
        if self.fused and named_grads and named_grads[0][1].is_cuda:
            if self._stepper is None:
                from grace_b200.parallel.fused import FusedStepper
                self._stepper = FusedStepper(self)
            outs = self._stepper.step_all(named_grads)
        if outs is None:
            outs = [self.step(g, n) for n, g in named_grads]
        if inplace and named_grads:
            torch._foreach_copy_([g for _, g in named_grads], [o.view_as(g) for o, (_, g) in zip(outs, named_grads)])
        return outs

    def close(self):
        """Release the native plans created by ``step`` / ``step_all`` (no-op on the generic path)."""
        if self._stepper is not None:
            self._stepper.close()


def _finish(compressor, decompressed, world_size):
    out = compressor.aggregate(decompressed)
    return out / world_size if compressor.average else out


class Allreduce(_NativeStep, Communicator):
    """Reduce in the compressed domain, then decompress once.  Only valid for linear,
    rank-invariantly shaped payloads (none, fp16, random-k values, PowerSGD 1-D)."""

    def __init__(self, compressor, memory, world_size, group=None):
        super().__init__(compressor, memory, world_size)
        self.group = group

    def send_receive(self, tensors, name, ctx):
        reduced = []
        for t in tensors:
            if not t.dtype.is_floating_point:
                raise TypeError(
                    "Allreduce needs a floating-point payload; %s sends %s (use Allgather)"
                    % (type(self.compressor).__name__, t.dtype))
            t = t.contiguous()
            if _live(self.world_size):
                dist.all_reduce(t, group=self.group)
            if self.compressor.average:
                t = t / self.world_size
            reduced.append(t)
        return self.compressor.decompress(reduced, ctx)


def gather_payload(tensors, world_size, same_size, group=None):
    """All-gather every payload tensor; returns ``per_rank[r] = [tensor, ...]``."""
    if not _live(world_size):
        return [list(tensors)]
    if same_size:
        columns = []
        for t in tensors:
            shape = t.shape
            w, restore = as_wire(t.contiguous())
            slots = [torch.empty_like(w) for _ in range(world_size)]
            dist.all_gather(slots, w, group=group)
            columns.append([restore(s).view(shape) for s in slots])
        return [list(col) for col in zip(*columns)]
    device = tensors[0].device
    mine = torch.tensor([t.numel() for t in tensors], dtype=torch.int64, device=device)
    sizes = [torch.empty_like(mine) for _ in range(world_size)]
    dist.all_gather(sizes, mine, group=group)
    sizes = torch.stack(sizes).tolist()                       # [rank][payload]
    columns = []
    for j, t in enumerate(tensors):
        flat = t.contiguous().view(-1)
        cap = max(s[j] for s in sizes)
        padded = flat.new_zeros(cap)
        padded[:flat.numel()] = flat
        w, restore = as_wire(padded)
        slots = [torch.empty_like(w) for _ in range(world_size)]
        dist.all_gather(slots, w, group=group)
        columns.append([restore(slot)[:sizes[r][j]] for r, slot in enumerate(slots)])
    return [list(col) for col in zip(*columns)]


class Allgather(_NativeStep, Communicator):
    """Gather every rank's payload, decompress each, aggregate."""

    def __init__(self, compressor, memory, world_size, group=None):
        super().__init__(compressor, memory, world_size)
        self.group = group

    def send_receive(self, tensors, name, ctx):
        per_rank = gather_payload(tensors, self.world_size, self.compressor.tensors_size_are_same, self.group)
        decompressed = [self.compressor.decompress(p, ctx) for p in per_rank]
        return _finish(self.compressor, decompressed, self.world_size)


class Broadcast(_NativeStep, Communicator):
    """Every rank broadcasts its payload in turn; same result as ``Allgather`` for
    fixed-size payloads."""

    def __init__(self, compressor, memory, world_size, rank=None, group=None):
        super().__init__(compressor, memory, world_size)
        if rank is None:
            rank = dist.get_rank(group) if (dist.is_available() and dist.is_initialized()) else 0
        self.rank = rank
        self.group = group

    def send_receive(self, tensors, name, ctx):
        if not self.compressor.tensors_size_are_same:
            raise NotImplementedError("Broadcast cannot carry variable-size payloads; use Allgather")
        decompressed = []
        for root in range(self.world_size if _live(self.world_size) else 1):
            received = []
            for t in tensors:
                w, restore = as_wire(t.contiguous())
                buf = w if root == self.rank else torch.empty_like(w)
                if _live(self.world_size):
                    dist.broadcast(buf, root, group=self.group)
                received.append(restore(buf).view(t.shape))
            decompressed.append(self.compressor.decompress(received, ctx))
        return _finish(self.compressor, decompressed, self.world_size)
