"""Horovod-style collective ops with integer handles
(API of ``patch_files/horovod/torch/mpi_ops.py:42-439`` in the reference tree, whose
native half is Horovod's C++ core ``mpi_lib_v2``).

No background negotiation thread is needed here: ops are issued in program order on every
rank, so each call maps directly onto an asynchronous ``torch.distributed`` op (plumbing /
CPU / Gloo) -- or, for gradient buckets owned by the fused engine, onto CUDA events created
by ``libgrace_b200.so`` (``grace_b200/parallel/engine.py``).  ``_handle_map`` keeps inputs
and outputs alive until ``synchronize`` like the reference does (``mpi_ops.py:60,88``).
"""
from __future__ import annotations

import os
import itertools

import torch
from torch import distributed as dist

from grace_b200.utils.wire import as_wire

_handle_map = {}
_counter = itertools.count(1)
_group = None


def init(backend=None, group=None, **kw):
    """Join the job.  Reads the usual ``RANK`` / ``WORLD_SIZE`` / ``MASTER_*`` variables;
    a single process without them runs as world size 1 with no process group."""
    global _group
    _group = group
    if dist.is_available() and not dist.is_initialized() and "RANK" in os.environ:
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if torch.cuda.is_available():
            torch.cuda.set_device(local_rank())
        dist.init_process_group(backend, **kw)


def shutdown():
    _handle_map.clear()
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


def _live():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size(_group) > 1


def size():
    return dist.get_world_size(_group) if (dist.is_available() and dist.is_initialized()) else 1


def rank():
    return dist.get_rank(_group) if (dist.is_available() and dist.is_initialized()) else 0


def local_rank():
    return int(os.environ.get("LOCAL_RANK", 0))


def local_size():
    return int(os.environ.get("LOCAL_WORLD_SIZE", 1))


def mpi_threads_supported():
    return False


def nccl_built():
    return dist.is_available() and dist.is_nccl_available()


def gloo_built():
    return dist.is_available() and dist.is_gloo_available()


class _Pending:
    __slots__ = ("works", "finish", "keep")

    def __init__(self, works, finish, keep):
        self.works, self.finish, self.keep = works, finish, keep


def _register(works, finish, keep=()):
    h = next(_counter)
    _handle_map[h] = _Pending([w for w in works if w is not None], finish, keep)
    return h


def poll(handle):
    """True once the op behind ``handle`` has finished (non-blocking)."""
    return all(w.is_completed() for w in _handle_map[handle].works)


def synchronize(handle):
    """Wait for ``handle``, release it, and return the op's output tensor."""
    if handle not in _handle_map:
        raise ValueError("unknown or already synchronized handle %r" % (handle,))
    pending = _handle_map.pop(handle)
    for w in pending.works:
        w.wait()
    return pending.finish()


# ---------------------------------------------------------------- allreduce
def allreduce_async_(tensor, average=True, name=None):
    work = dist.all_reduce(tensor, group=_group, async_op=True) if _live() else None
    n = size()

    def finish():
        if average and n > 1:
            tensor.div_(n)
        return tensor
    return _register([work], finish, (tensor,))


def allreduce_async(tensor, average=True, name=None):
    return allreduce_async_(tensor.clone(), average, name)


def allreduce_(tensor, average=True, name=None):
    return synchronize(allreduce_async_(tensor, average, name))


class _AllreduceFn(torch.autograd.Function):
    """Differentiable all-reduce (reference: ``HorovodAllreduce``, patch_files/horovod/torch/mpi_ops.py:116-127):
    the gradient of a sum / mean over ranks is the same reduction of the upstream gradients."""

    @staticmethod
    def forward(ctx, tensor, average, name):
        ctx.average = average
        return synchronize(allreduce_async(tensor.detach(), average, name))

    @staticmethod
    def backward(ctx, grad_output):
        return synchronize(allreduce_async(grad_output.contiguous(), ctx.average, None)), None, None


def allreduce(tensor, average=True, name=None):
    if tensor.requires_grad and torch.is_grad_enabled():
        return _AllreduceFn.apply(tensor, average, name)
    return synchronize(allreduce_async(tensor, average, name))


# ---------------------------------------------------------------- allgather
def allgather_async(tensor, name=None):
    """Concatenate along dim 0; first dimensions may differ between ranks."""
    if tensor.dim() == 0:
        tensor = tensor.reshape(1)
    if not _live():
        return _register([], lambda: tensor, (tensor,))
    n = size()
    mine = torch.tensor([tensor.shape[0]], dtype=torch.int64, device=tensor.device)
    sizes = [torch.empty_like(mine) for _ in range(n)]
    dist.all_gather(sizes, mine, group=_group)
    sizes = [int(s) for s in sizes]
    cap = max(sizes)
    padded = tensor.new_zeros((cap,) + tuple(tensor.shape[1:]))
    padded[:tensor.shape[0]] = tensor
    wire, restore = as_wire(padded)
    slots = [torch.empty_like(wire) for _ in range(n)]
    work = dist.all_gather(slots, wire, group=_group, async_op=True)
    return _register([work], lambda: torch.cat([restore(s).view(padded.shape)[:k] for s, k in zip(slots, sizes)],
                                               dim=0), (wire, slots))


class _AllgatherFn(torch.autograd.Function):
    """Differentiable all-gather (reference: ``HorovodAllgather``, mpi_ops.py:237-255): the upstream gradient is
    summed over ranks and every rank keeps the rows it contributed."""

    @staticmethod
    def forward(ctx, tensor, name):
        ctx.rows = tensor.shape[0] if tensor.dim() else 1
        ctx.scalar = tensor.dim() == 0
        rows = synchronize(allgather_async(torch.tensor([ctx.rows], dtype=torch.int64, device=tensor.device), None))
        ctx.offset = int(rows[:rank()].sum())
        return synchronize(allgather_async(tensor.detach(), name))

    @staticmethod
    def backward(ctx, grad_output):
        total = synchronize(allreduce_async(grad_output.contiguous(), False, None))
        mine = total.narrow(0, ctx.offset, ctx.rows)
        return (mine.reshape(()) if ctx.scalar else mine), None


def allgather(tensor, name=None):
    if tensor.requires_grad and torch.is_grad_enabled():
        return _AllgatherFn.apply(tensor, name)
    return synchronize(allgather_async(tensor, name))


# ---------------------------------------------------------------- broadcast
def broadcast_async_(tensor, root_rank, name=None):
    if not _live():
        return _register([], lambda: tensor, (tensor,))
    wire, restore = as_wire(tensor)
    work = dist.broadcast(wire, root_rank, group=_group, async_op=True)

    def finish():
        if wire is not tensor:
            tensor.copy_(restore(wire).view(tensor.shape))
        return tensor
    return _register([work], finish, (tensor, wire))


def broadcast_async(tensor, root_rank, name=None):
    return broadcast_async_(tensor.clone(), root_rank, name)


def broadcast_(tensor, root_rank, name=None):
    return synchronize(broadcast_async_(tensor, root_rank, name))


class _BroadcastFn(torch.autograd.Function):
    """Differentiable broadcast (reference: ``HorovodBroadcast``, mpi_ops.py:319-333): the root receives the sum of
    all ranks' upstream gradients, every other rank a zero gradient."""

    @staticmethod
    def forward(ctx, tensor, root_rank, name):
        ctx.root = root_rank
        return synchronize(broadcast_async(tensor.detach(), root_rank, name))

    @staticmethod
    def backward(ctx, grad_output):
        total = synchronize(allreduce_async(grad_output.contiguous(), False, None))
        if rank() != ctx.root:
            total = torch.zeros_like(total)
        return total, None, None


def broadcast(tensor, root_rank, name=None):
    if tensor.requires_grad and torch.is_grad_enabled():
        return _BroadcastFn.apply(tensor, root_rank, name)
    return synchronize(broadcast_async(tensor, root_rank, name))
