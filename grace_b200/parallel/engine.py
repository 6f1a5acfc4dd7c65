"""Communicator / plan objects over the native library.

``Engine`` is what Horovod's ``hvd.init()`` + C++ core is to the reference
(``patch_files/horovod/torch/mpi_ops.py:42-55``): rank bootstrap, the peer-memory arena and
the handle table.  The IPC handles travel once through ``torch.distributed``
(``all_gather_object``); after that no collective library is involved in gradient exchange.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Sequence

import torch
from torch import distributed as dist

from grace_b200.ops import native
from grace_b200.ops.native import CodecParams, check
from grace_b200.utils.philox import fnv1a


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


class Engine:
    """One rank's communicator.  ``Engine.create()`` joins the torch.distributed job (or runs
    as a single rank); ``Engine.local(world)`` builds ``world`` virtual ranks on the current
    GPU for protocol tests on one device."""

    def __init__(self, handle, rank, world, group=None):
        self._h = C.c_void_p(handle)
        self.rank, self.world, self.group = rank, world, group
        self._local = False
        self._symmetric = []
        self.multicast = False
        self.nvls_error = None
        self.plans: List["Plan"] = []

    # ------------------------------------------------------------------ construction
    @staticmethod
    def default_arena_bytes():
        return int(os.environ.get("GRACE_B200_ARENA_MB", "1024")) << 20

    @classmethod
    def create(cls, arena_bytes=None, group=None):
        L = native.lib()
        if not torch.cuda.is_available():
            raise native.NativeError("the fused engine needs a CUDA device")
        check(L.gb_set_device(torch.cuda.current_device()))
        live = dist.is_available() and dist.is_initialized()
        rank = dist.get_rank(group) if live else 0
        world = dist.get_world_size(group) if live else 1
        h = C.c_void_p()
        check(L.gb_comm_create(rank, world, arena_bytes or cls.default_arena_bytes(), C.byref(h)))
        eng = cls(h.value, rank, world, group)
        if world > 1:
            buf = C.create_string_buffer(64)
            check(L.gb_comm_ipc_handle(eng._h, buf))
            gathered = [None] * world
            dist.all_gather_object(gathered, bytes(buf.raw), group=group)
            check(L.gb_comm_open_peers(eng._h, b"".join(gathered)))
            dist.barrier(group=group)
            if os.environ.get("GRACE_B200_NVLS_PAYLOAD", "1") != "0" and group is None:
                eng._adopt_symmetric(arena_bytes or cls.default_arena_bytes())
        return eng

    @classmethod
    def local(cls, world, arena_bytes=None):
        L = native.lib()
        check(L.gb_set_device(torch.cuda.current_device()))
        arr = (C.c_void_p * world)()
        check(L.gb_comm_create_local(world, arena_bytes or (256 << 20), arr))
        engines = [cls(arr[r], r, world) for r in range(world)]
        for e in engines:
            e._local = True
        return engines

    # ------------------------------------------------------------------ services
    def status(self):
        out = C.c_uint32()
        check(native.lib().gb_comm_status(self._h, C.byref(out)))
        return out.value

    def check_status(self):
        if os.environ.get("GRACE_B200_GUARD", "0") not in ("", "0") and not torch.cuda.is_current_stream_capturing():
            check(native.lib().gb_comm_check_guards(self._h, None))      # debugging runs: scan the guard bands first
        s = self.status()
        if s:
            native.lib().gb_comm_clear_status(self._h)
            what = [n for b, n in ((1, "peer flag wait timed out (a rank is missing or hung)"),
                                   (2, "variable-size payload exceeded the slot capacity"),
                                   (4, "a guard band behind a wire slot was overwritten"),
                                   (8, "a bounds assertion of the debug build failed")) if s & b]
            raise native.NativeError("engine status 0x%x: %s" % (s, "; ".join(what)))

    def check_guards(self):
        """Scan the poisoned guard bands behind this rank's wire slots (``GRACE_B200_GUARD=1``); raises through
        ``check_status`` when a peer store ran past its slot.  Returns the number of guarded channels."""
        n = C.c_int()
        check(native.lib().gb_comm_check_guards(self._h, C.byref(n)))
        self.check_status()
        return n.value

    def set_timeout_ms(self, ms):
        check(native.lib().gb_comm_set_timeout_ms(self._h, int(ms)))

    def arena_free(self):
        return native.lib().gb_comm_arena_free(self._h)

    def _adopt_symmetric(self, nbytes):
        """Make the next arena region a symmetric allocation bound to an NVLS multicast address (when the host has
        NVLink SHARP): wire slots carved from it are written with ONE multimem.st per 16 bytes for all ranks -- the
        switch replicates the store -- instead of W - 1 peer stores.  Collective; all ranks agree or nobody adopts
        (agreed fallback: the CUDA IPC arena with P2P store loops)."""
        import torch.distributed._symmetric_memory as symm
        L = native.lib()
        buf = handle = None
        ok = 1
        try:
            buf = symm.empty(int(nbytes), dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
            handle = symm.rendezvous(buf, dist.group.WORLD)
            mc = int(handle.multicast_ptr or 0)
            ptrs = [int(x) for x in handle.buffer_ptrs]
            if not mc or len(ptrs) != self.world:
                ok = 0
        except Exception as exc:                              # no multicast on this host / torch without symmetric memory
            self.nvls_error = repr(exc)
            ok = 0
        t = torch.tensor([ok], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        if int(t) == 0:
            return False
        buf.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        arr = (C.c_uint64 * self.world)(*ptrs)
        idx = C.c_int()
        check(L.gb_comm_region_adopt(self._h, arr, C.c_uint64(mc), int(nbytes), C.byref(idx)))
        self._symmetric.append((buf, handle))                 # the memory must outlive the communicator
        self.multicast = True
        return True

    def grow(self, nbytes):
        """Collective: every rank adds an arena region of at least ``nbytes`` (same size everywhere), exports it
        with CUDA IPC and maps the peers' regions.  Called automatically when a plan's wire slot does not fit,
        so the arena is not limited to what the first model needed."""
        L = native.lib()
        nbytes = max(int(nbytes) + (1 << 20), int(os.environ.get("GRACE_B200_ARENA_GROW_MB", "256")) << 20)
        if self.multicast and self._adopt_symmetric(nbytes):
            return native.lib().gb_comm_regions(self._h) - 1
        idx = C.c_int()
        check(L.gb_comm_region_add(self._h, nbytes, C.byref(idx)))
        if self.world > 1 and not self._local:
            buf = C.create_string_buffer(64)
            check(L.gb_comm_region_ipc_handle(self._h, idx.value, buf))
            gathered = [None] * self.world
            dist.all_gather_object(gathered, bytes(buf.raw), group=self.group)
            check(L.gb_comm_region_open_peers(self._h, idx.value, b"".join(gathered)))
            dist.barrier(group=self.group)
        return idx.value

    def create_with_growth(self, create):
        """Run ``create()`` (a native plan constructor returning its status); on "arena exhausted" grow the
        arena by what was missing and retry."""
        L = native.lib()
        rc = -1
        for _ in range(4):
            rc = create()
            if rc == 0:
                return
            need = L.gb_comm_arena_need(self._h)
            if not need:
                break
            self.grow(need)
        check(rc)

    def barrier(self, stream=None):
        check(native.lib().gb_comm_barrier(self._h, _stream_ptr(stream)))

    def destroy(self):
        for p in list(self.plans):
            p.destroy()
        if self._h:
            native.lib().gb_comm_destroy(self._h)
            self._h = None

    def plan(self, numels: Sequence[int], names: Sequence[str], params: CodecParams) -> "Plan":
        return Plan(self, numels, names, params)


class Plan:
    """One gradient bucket: geometry + error-feedback state + wire slot (``gb_plan``)."""

    def __init__(self, engine: Engine, numels, names, params: CodecParams):
        L = native.lib()
        self.engine = engine
        self.nseg = len(numels)
        self.numels = [int(n) for n in numels]
        self.names = list(names)
        self.params = params
        nm = (C.c_int64 * self.nseg)(*self.numels)
        hs = (C.c_uint32 * self.nseg)(*[fnv1a(n) for n in names])
        h = C.c_void_p()
        engine.create_with_growth(lambda: L.gb_plan_create(engine._h, self.nseg, nm, hs, C.byref(params), C.byref(h)))
        self._h = h
        self.padded = L.gb_plan_padded_numel(h)
        off = (C.c_int64 * self.nseg)()
        check(L.gb_plan_offsets(h, off))
        self.offsets = list(off)
        self.wire_bytes = L.gb_plan_wire_bytes(h)
        engine.plans.append(self)

    def send(self, grad: torch.Tensor, stream=None):
        assert grad.is_cuda and grad.dtype == torch.float32 and grad.numel() >= self.padded
        check(native.lib().gb_plan_send(self._h, C.c_void_p(grad.data_ptr()), _stream_ptr(stream)))

    def presend(self, grad: torch.Tensor, stream=None):
        """First half of ``send`` for plans whose send waits for the peers (DGC gradient clipping); only needed when
        several ranks are driven from one stream (``Engine.local``)."""
        check(native.lib().gb_plan_presend(self._h, C.c_void_p(grad.data_ptr()), _stream_ptr(stream)))

    def mid(self, stream=None):
        check(native.lib().gb_plan_mid(self._h, _stream_ptr(stream)))

    def wait(self, stream=None):
        """One-warp wait for the peers' flags of the exchange in flight (keeps the receive grids off the SMs)."""
        check(native.lib().gb_plan_wait(self._h, _stream_ptr(stream)))

    def recv(self, out: torch.Tensor, stream=None):
        assert out.is_cuda and out.dtype == torch.float32 and out.numel() >= self.padded
        check(native.lib().gb_plan_recv(self._h, C.c_void_p(out.data_ptr()), _stream_ptr(stream)))

    def launches(self):
        out = C.c_int64()
        check(native.lib().gb_plan_launches(self._h, C.byref(out)))
        return out.value

    # ---- state (checkpointing): copies, so a state_dict outlives the plan
    def _state_view(self, which):
        ptr, size = C.c_void_p(), C.c_size_t()
        check(native.lib().gb_plan_state(self._h, which, C.byref(ptr), C.byref(size)))
        return ptr.value, size.value

    def get_state(self, which):
        """Copy of the plan-owned state buffer (padded, float32) or ``None``."""
        ptr, size = self._state_view(which)
        if not ptr:
            return None
        out = torch.empty(size // 4, dtype=torch.float32, device="cuda")
        check(native.lib().gb_plan_state_read(self._h, which, C.c_void_p(out.data_ptr()), _stream_ptr()))
        return out

    def set_state(self, which, tensor):
        ptr, size = self._state_view(which)
        if not ptr:
            return
        src = tensor.to(device="cuda", dtype=torch.float32).reshape(-1).contiguous()
        assert src.numel() * 4 == size, "state size mismatch"
        check(native.lib().gb_plan_state_write(self._h, which, C.c_void_p(src.data_ptr()), _stream_ptr()))
        torch.cuda.current_stream().synchronize()

    @property
    def step(self):
        out = C.c_uint32()
        check(native.lib().gb_plan_get_step(self._h, C.byref(out)))
        return out.value

    @step.setter
    def step(self, value):
        check(native.lib().gb_plan_set_step(self._h, int(value)))

    def destroy(self):
        if self._h:
            native.lib().gb_plan_destroy(self._h)
            self._h = None
            if self in self.engine.plans:
                self.engine.plans.remove(self)


class PowerSGDPlan:
    """All >=2-D gradients of one bucket through the native PowerSGD pipeline (``gb_psgd``)."""

    def __init__(self, engine: Engine, shapes, names, rank, seed, has_memory, use_tcgen05=True, warm_start=False):
        L = native.lib()
        self.engine = engine
        self.nmat = len(shapes)
        self.shapes = [(int(r), int(c)) for r, c in shapes]
        self.numels = [r * c for r, c in self.shapes]
        self.names = list(names)
        rows = (C.c_int64 * self.nmat)(*[r for r, _ in self.shapes])
        cols = (C.c_int64 * self.nmat)(*[c for _, c in self.shapes])
        hs = (C.c_uint32 * self.nmat)(*[fnv1a(n) for n in names])
        h = C.c_void_p()
        engine.create_with_growth(lambda: L.gb_psgd_create2(engine._h, self.nmat, rows, cols, hs, int(rank), int(seed) & 0xFFFFFFFF,
                                                            1 if has_memory else 0, 1 if use_tcgen05 else 0,
                                                            1 if warm_start else 0, C.byref(h)))
        self.warm_start = bool(warm_start)
        self._h = h
        self.padded = L.gb_psgd_padded_numel(h)
        off = (C.c_int64 * self.nmat)()
        check(L.gb_psgd_offsets(h, off))
        self.offsets = list(off)
        self.wire_bytes = L.gb_psgd_wire_bytes(h)
        self.tc_matrices = L.gb_psgd_tc_matrices(h)
        engine.plans.append(self)

    def send(self, grad, stream=None):
        assert grad.is_cuda and grad.dtype == torch.float32 and grad.numel() >= self.padded
        check(native.lib().gb_psgd_send(self._h, C.c_void_p(grad.data_ptr()), _stream_ptr(stream)))

    def mid(self, stream=None):
        check(native.lib().gb_psgd_mid(self._h, _stream_ptr(stream)))

    def wait(self, which=0, stream=None):
        check(native.lib().gb_psgd_wait(self._h, int(which), _stream_ptr(stream)))

    def recv(self, out, stream=None):
        assert out.is_cuda and out.dtype == torch.float32 and out.numel() >= self.padded
        check(native.lib().gb_psgd_recv(self._h, C.c_void_p(out.data_ptr()), _stream_ptr(stream)))

    def launches(self):
        out = C.c_int64()
        check(native.lib().gb_psgd_launches(self._h, C.byref(out)))
        return out.value

    def get_state(self, which=0):
        """0: residual (error feedback), 1: the previous exchange's Q of a warm-started plan."""
        ptr, size = C.c_void_p(), C.c_size_t()
        if which == 1:
            check(native.lib().gb_psgd_qprev(self._h, C.byref(ptr), C.byref(size)))
            if not ptr.value:
                return None
            out = torch.empty(size.value // 4, dtype=torch.float32, device="cuda")
            check(native.lib().gb_psgd_qprev_copy(self._h, C.c_void_p(out.data_ptr()), 0, _stream_ptr()))
            return out
        check(native.lib().gb_psgd_residual(self._h, C.byref(ptr), C.byref(size)))
        if which != 0 or not ptr.value:
            return None
        out = torch.empty(size.value // 4, dtype=torch.float32, device="cuda")
        check(native.lib().gb_psgd_state_read(self._h, C.c_void_p(out.data_ptr()), _stream_ptr()))
        return out

    def set_state(self, which, tensor):
        if which == 1 and tensor is not None and self.warm_start:
            src = tensor.to(device="cuda", dtype=torch.float32).reshape(-1).contiguous()
            check(native.lib().gb_psgd_qprev_copy(self._h, C.c_void_p(src.data_ptr()), 1, _stream_ptr()))
            torch.cuda.current_stream().synchronize()
            return
        if which != 0 or tensor is None:
            return
        src = tensor.to(device="cuda", dtype=torch.float32).reshape(-1).contiguous()
        check(native.lib().gb_psgd_state_write(self._h, C.c_void_p(src.data_ptr()), _stream_ptr()))
        torch.cuda.current_stream().synchronize()

    @property
    def step(self):
        out = C.c_uint32()
        check(native.lib().gb_psgd_get_step(self._h, C.byref(out)))
        return out.value

    @step.setter
    def step(self, value):
        check(native.lib().gb_psgd_set_step(self._h, int(value)))

    def destroy(self):
        if self._h:
            native.lib().gb_psgd_destroy(self._h)
            self._h = None
            if self in self.engine.plans:
                self.engine.plans.remove(self)


class NvlsPlan:
    """In-place NVLS all-reduce of one bucket that lives in multicast-bound symmetric memory."""

    def __init__(self, engine: Engine, padded: int, multicast_ptr: int, average=True):
        h = C.c_void_p()
        check(native.lib().gb_nvls_create(engine._h, int(padded), C.byref(h)))
        self._h, self.engine, self.padded = h, engine, int(padded)
        self.mc, self.average = int(multicast_ptr), bool(average)
        self.wire_bytes = (self.padded * 4) // max(engine.world, 1)
        self.step = 0
        engine.plans.append(self)

    def signal(self, stream=None):
        check(native.lib().gb_nvls_signal(self._h, _stream_ptr(stream)))

    def wait(self, stream=None):
        pass

    def reduce(self, stream=None):
        check(native.lib().gb_nvls_reduce(self._h, C.c_void_p(self.mc), 1 if self.average else 0, _stream_ptr(stream)))
        self.step += 1

    def launches(self):
        out = C.c_int64()
        check(native.lib().gb_nvls_launches(self._h, C.byref(out)))
        return out.value

    def get_state(self, which):
        return None

    def set_state(self, which, tensor):
        pass

    def destroy(self):
        if self._h:
            native.lib().gb_nvls_destroy(self._h)
            self._h = None
            if self in self.engine.plans:
                self.engine.plans.remove(self)
