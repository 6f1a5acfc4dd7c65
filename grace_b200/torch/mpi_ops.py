"""Horovod-style collective ops with integer handles
(API of ``patch_files/horovod/torch/mpi_ops.py:42-439`` in the reference tree, whose
native half is Horovod's C++ core ``mpi_lib_v2``).

No background negotiation thread is needed here: ops are issued in program order on every
rank.  CUDA tensors of a job that lives on one NVLink domain take the NATIVE path: the op runs
as two small kernels over the engine's peer arena (``csrc/ops.cu``: P2P stores + epoch flags;
large fp32 all-reduces through the dense two-shot plans of ``csrc/k0_dense.cu``) on a dedicated
stream, and the integer handle is an event handle of ``libgrace_b200.so`` (``gb_handle_record``
/ ``gb_handle_poll`` / ``gb_handle_wait``) -- no collective library is involved.  Everything
else (CPU tensors, Gloo, process sub-groups, jobs spanning hosts) maps onto an asynchronous
``torch.distributed`` op.  ``_handle_map`` keeps inputs and outputs alive until ``synchronize``
like the reference does (``mpi_ops.py:60,88``).  ``GRACE_B200_NATIVE_OPS=0`` forces the generic path.
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


def native_op_launches():
    """Kernels launched so far by the native op layer (0 when every op went through torch.distributed)."""
    return _native.ops_launched + sum(pl.launches() for pl, _ in _native.dense_plans.values())


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
    __slots__ = ("works", "finish", "keep", "event")

    def __init__(self, works, finish, keep, event=None):
        self.works, self.finish, self.keep, self.event = works, finish, keep, event


def _register(works, finish, keep=(), event=None):
    h = next(_counter)
    _handle_map[h] = _Pending([w for w in works if w is not None], finish, keep, event)
    return h


def poll(handle):
    """True once the op behind ``handle`` has finished (non-blocking)."""
    pending = _handle_map[handle]
    if pending.event is not None:                       # native op: event handle of libgrace_b200
        from grace_b200.ops import native
        rc = native.lib().gb_handle_poll(pending.event)
        if rc < 0:
            native.check(rc)
        return bool(rc)
    return all(w.is_completed() for w in pending.works)


def synchronize(handle):
    """Wait for ``handle``, release it, and return the op's output tensor."""
    if handle not in _handle_map:
        raise ValueError("unknown or already synchronized handle %r" % (handle,))
    pending = _handle_map.pop(handle)
    if pending.event is not None:
        from grace_b200.ops import native
        native.check(native.lib().gb_handle_wait(pending.event))       # blocks the host like horovod_torch_wait_and_clear
        _native.engine.check_status()
    for w in pending.works:
        w.wait()
    return pending.finish()


# ---------------------------------------------------------------- native path (csrc/ops.cu)
class _NativeOps:
    """Point collectives over the engine's arena, issued in program order on one dedicated stream."""

    SMALL_ALLREDUCE = 4 << 20            # bytes: up to here an all-reduce is "every rank pushes, every rank sums"
    _MODES = {torch.float32: 1, torch.float64: 2, torch.int64: 3, torch.int32: 4}

    def __init__(self):
        self.engine = None
        self.ops = None
        self.stream = None
        self.dense_plans = {}
        self.enabled = None
        self.ops_launched = 0

    def usable(self, tensor):
        if self.enabled is None:
            ok = (os.environ.get("GRACE_B200_NATIVE_OPS", "1") != "0" and _live() and _group is None
                  and torch.cuda.is_available())
            if ok:
                from grace_b200.ops import native
                from grace_b200.parallel.fused import single_nvlink_domain
                ok = native.available() and single_nvlink_domain()
            self.enabled = ok
        return bool(self.enabled and tensor.is_cuda and tensor.numel() > 0
                    and tensor.device.index == torch.cuda.current_device())

    def _setup(self, nbytes):
        import ctypes as C
        from grace_b200.ops import native
        from grace_b200.parallel.fused import get_engine
        if self.engine is None:
            self.engine = get_engine()
            self.stream = torch.cuda.Stream()
        L = native.lib()
        if self.ops is None or L.gb_ops_capacity(self.ops) < nbytes:
            cap = 1 << 20
            while cap < nbytes:
                cap <<= 1
            h = C.c_void_p()
            self.engine.create_with_growth(lambda: L.gb_ops_create(self.engine._h, cap, C.byref(h)))
            self.ops = h                                  # an outgrown channel stays allocated (the arena only grows)
        return L

    def _enter(self):
        ready = torch.cuda.Event()
        ready.record()
        self.stream.wait_event(ready)

    def _handle(self):
        import ctypes as C
        from grace_b200.ops import native
        h = C.c_int64()
        native.check(native.lib().gb_handle_record(C.c_void_p(self.stream.cuda_stream), C.byref(h)))
        return h.value

    @staticmethod
    def _staging(tensor):
        """Contiguous bytes of ``tensor`` padded to a multiple of 16 (own buffer: the op runs on another stream)."""
        raw = tensor.detach().contiguous().view(-1).view(torch.uint8)
        n = raw.numel()
        buf = torch.zeros((n + 15) // 16 * 16, dtype=torch.uint8, device=tensor.device)
        buf[:n].copy_(raw)
        return buf, n

    def exchange(self, src, nbytes, first, count, stride, mode, scale, out):
        import ctypes as C
        from grace_b200.ops import native
        L = self._setup(nbytes)
        sp = C.c_void_p(self.stream.cuda_stream)
        native.check(L.gb_ops_send(self.ops, C.c_void_p(src.data_ptr() if src is not None else 0), nbytes, sp))
        native.check(L.gb_ops_recv(self.ops, nbytes, first, count, stride, mode, C.c_double(scale),
                                   C.c_void_p(out.data_ptr()), sp))
        self.ops_launched += 2

    # -- the three ops; each returns (event handle, finish, keep-alive)
    def allreduce_(self, tensor, average):
        n = size()
        work = tensor if tensor.dtype in self._MODES else tensor.float()
        nbytes = work.numel() * work.element_size()
        if nbytes > self.SMALL_ALLREDUCE:
            if work.dtype != torch.float32:
                return None
            return self._dense_allreduce(tensor, work, average)
        buf, nb = self._staging(work)
        self._setup(buf.numel())
        self._enter()
        out = torch.empty_like(buf)
        scale = 1.0 / n if (average and work.dtype.is_floating_point) else 1.0
        with torch.cuda.stream(self.stream):
            self.exchange(buf, buf.numel(), 0, n, buf.numel(), self._MODES[work.dtype], scale, out)
        event = self._handle()

        def finish():
            res = out[:nb].view(work.dtype).view(tensor.shape)
            if average and not work.dtype.is_floating_point:
                res = torch.div(res, n, rounding_mode="floor") if not tensor.dtype.is_floating_point else res / n
            tensor.copy_(res)
            return tensor
        return event, finish, (tensor, buf, out)

    def _dense_allreduce(self, tensor, work, average):
        from grace_b200.ops.native import CodecParams, CODEC, MEMORY
        self._setup(16)
        key = (work.numel(), bool(average))
        entry = self.dense_plans.get(key)
        if entry is None:
            prm = CodecParams()
            prm.codec, prm.memory, prm.average, prm.beta, prm.gamma = CODEC["none"], MEMORY["none"], 1 if average else 0, 1.0, 1.0
            plan = self.engine.plan([work.numel()], ["hvd.allreduce.%d" % work.numel()], prm)
            entry = (plan, torch.zeros(plan.padded, dtype=torch.float32, device=tensor.device))
            self.dense_plans[key] = entry
        plan, flat = entry
        self._enter()
        with torch.cuda.stream(self.stream):
            flat[:work.numel()].copy_(work.reshape(-1))
            plan.send(flat, stream=self.stream)
            plan.mid(stream=self.stream)
            plan.recv(flat, stream=self.stream)
            tensor.copy_(flat[:work.numel()].view(tensor.shape))
        return self._handle(), (lambda: tensor), (tensor, work)

    def allgather(self, tensor):
        n = size()
        rows = torch.zeros(2, dtype=torch.int64, device=tensor.device)
        rows[0] = tensor.shape[0]
        sizes_out = torch.empty(2 * n, dtype=torch.int64, device=tensor.device)
        self._setup(16)
        self._enter()
        with torch.cuda.stream(self.stream):
            self.exchange(rows, 16, 0, n, 16, 0, 1.0, sizes_out)
        self.stream.synchronize()                               # the sizes are needed on the host (the reference blocks too)
        sizes = [int(v) for v in sizes_out.view(n, 2)[:, 0].tolist()]
        cap = max(sizes)
        row_bytes = int(torch.Size(tensor.shape[1:]).numel()) * tensor.element_size()
        padded = torch.zeros((cap,) + tuple(tensor.shape[1:]), dtype=tensor.dtype, device=tensor.device)
        padded[:tensor.shape[0]] = tensor
        buf, nb = self._staging(padded)
        out = torch.empty(n * buf.numel(), dtype=torch.uint8, device=tensor.device)
        self._enter()
        with torch.cuda.stream(self.stream):
            self.exchange(buf, buf.numel(), 0, n, buf.numel(), 0, 1.0, out)
        event = self._handle()

        def finish():
            parts = [out[r * buf.numel():r * buf.numel() + sizes[r] * row_bytes].view(tensor.dtype)
                     .view((sizes[r],) + tuple(tensor.shape[1:])) for r in range(n)]
            return torch.cat(parts, dim=0)
        return event, finish, (tensor, buf, out)

    def broadcast_(self, tensor, root_rank):
        buf, nb = self._staging(tensor)
        out = torch.empty_like(buf)
        self._setup(buf.numel())
        self._enter()
        with torch.cuda.stream(self.stream):
            self.exchange(buf if rank() == root_rank else None, buf.numel(), root_rank, 1, buf.numel(), 0, 1.0, out)
        event = self._handle()

        def finish():
            if rank() != root_rank:
                tensor.copy_(out[:nb].view(tensor.dtype).view(tensor.shape))
            return tensor
        return event, finish, (tensor, buf, out)


_native = _NativeOps()


def _try_native(fn, *args):
    got = fn(*args)
    if got is None:
        return None
    event, finish, keep = got
    return _register([], finish, keep, event=event)


# ---------------------------------------------------------------- allreduce
def allreduce_async_(tensor, average=True, name=None):
    if _native.usable(tensor):
        h = _try_native(_native.allreduce_, tensor, average)
        if h is not None:
            return h
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
    if _native.usable(tensor):
        return _try_native(_native.allgather, tensor)
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
    if _native.usable(tensor):
        return _try_native(_native.broadcast_, tensor, root_rank)
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
