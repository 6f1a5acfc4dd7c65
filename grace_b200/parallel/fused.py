"""Fused data-parallel gradient exchange: bucketing + overlap + dispatch on built-in triples.

Where the reference runs, per parameter, ``compensate -> compress -> update -> collective ->
W x decompress -> aggregate`` as ~12 + 3W library kernels and one NCCL/Horovod collective per
payload tensor (``grace_dl/dist/__init__.py:47-51``, ``communicator/allgather.py:8-45``;
SURVEY §3.2-3.3), this module maps every built-in ``(memory, compressor, communicator)``
triple onto native *plans* (``csrc/plan.cu``): one send launch sequence and one receive kernel
per ~25 MB gradient bucket, with error-feedback state in flat device buffers.

* ``FusedDP``      -- Horovod-style: gradients are views into flat buckets; a bucket's send
  is enqueued on a side stream the moment autograd finishes its last gradient (overlap with
  the rest of backward); receives run in ``synchronize()``.
* ``FusedStepper`` -- synchronous ``Communicator.step(tensor, name)`` fast path: one
  single-segment plan per tensor name, created on first use.

User-defined subclasses of the ABCs never reach this module (dispatch is on exact types) and
run through the generic torch.distributed path instead.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import torch

from grace_b200 import compressor as C
from grace_b200 import memory as M

DEFAULT_BUCKET_BYTES = int(os.environ.get("GRACE_B200_BUCKET_MB", "32")) << 20

_engine = None


def get_engine(min_arena_bytes=0):
    """Process-wide communicator (created on first use, like ``hvd.init()``).  The peer arena is
    sized for the first caller's needs (``min_arena_bytes``) or ``GRACE_B200_ARENA_MB``, whichever
    is larger; later plans that do not fit raise a clear error from the native allocator."""
    global _engine
    if _engine is None:
        from grace_b200.parallel.engine import Engine
        _engine = Engine.create(arena_bytes=max(Engine.default_arena_bytes(), int(min_arena_bytes)))
    return _engine


_topology_ok = None


def single_nvlink_domain():
    """The peer-memory engine needs every rank on one host (CUDA IPC + NVLink/NVSwitch) and at most
    8 ranks.  Jobs that span hosts keep working through the generic torch.distributed path."""
    global _topology_ok
    if _topology_ok is None:
        import socket
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            _topology_ok = True
        elif dist.get_world_size() > 8:
            _topology_ok = False
        else:
            hosts = [None] * dist.get_world_size()
            dist.all_gather_object(hosts, socket.gethostname())
            _topology_ok = len(set(hosts)) == 1
    return _topology_ok


def estimate_arena_bytes(numels, params, powersgd, world):
    """Upper bound of the peer-arena bytes the plans of one model need (2 parities x W slots)."""
    n = sum((x + 127) // 128 * 128 for x in numels)
    if powersgd is not None:
        # P and Q tiles: (rows + cols) x 4 floats per matrix, bounded by numel / 2 floats in total;
        # plus the dense plan of the 1-D tensors (bounded by the whole bucket)
        slot = n * 2 + (1 << 20) + (4 * n if world == 1 else 8 * n // world + 4096)
    else:
        from grace_b200.ops.native import CODEC
        code = params.codec
        if code in (CODEC["threshold"], CODEC["dgc"], CODEC["adaq"]):
            slot = int(8 * n * (params.capacity or 1.0))
        elif code == CODEC["inceptionn"]:
            slot = int(4 * n * (params.capacity or 1.0)) + n // 4 + 272 * (n // 4096 + len(numels))   # tags + per-tile headers
        elif code in (CODEC["topk"], CODEC["randomk"]):
            slot = int(8 * (n * params.ratio + len(numels)))
        elif code in (CODEC["none"], CODEC["fp16"]):
            slot = 4 * n if world == 1 else 8 * n // world + 4096
        else:
            slot = 2 * n                          # <= 16 bits per element
        slot += 16 * len(numels) + 8 * (n // 4096 + len(numels)) + 4096
    return int(slot * 2 * world * 1.1) + (64 << 20)


def shutdown_engine():
    global _engine
    if _engine is not None:
        _engine.destroy()
        _engine = None


def codec_params(compressor, memory):
    """Translate a built-in (compressor, memory) pair into ``gb_codec_params``; ``None`` if the
    pair has no native implementation (user classes; PowerSGD has its own plan type)."""
    from grace_b200.ops.native import CodecParams, CODEC, MEMORY
    p = CodecParams()
    p.beta = p.gamma = 1.0
    p.capacity = float(os.environ.get("GRACE_B200_SPARSE_CAPACITY", "1.0"))
    p.average = 1 if compressor.average else 0
    p.seed = getattr(compressor, "seed", 0)
    t = type(compressor)
    if t is C.NoneCompressor:
        p.codec = CODEC["none"]
    elif t is C.FP16Compressor and compressor.wire_dtype == torch.float16:
        p.codec = CODEC["fp16"]
    elif t is C.TopKCompressor:
        p.codec, p.ratio = CODEC["topk"], compressor.compress_ratio
    elif t is C.RandomKCompressor:
        p.codec, p.ratio = CODEC["randomk"], compressor.compress_ratio
    elif t is C.ThresholdCompressor:
        p.codec, p.threshold = CODEC["threshold"], compressor.threshold
    elif t is C.DgcCompressor:
        p.codec, p.ratio = CODEC["dgc"], compressor.compress_ratio
    elif t is C.QSGDCompressor:
        p.codec, p.quantum_num = CODEC["qsgd"], int(compressor.quantum_num)
        if compressor.quantum_num >= 32768:
            return None
    elif t is C.TernGradCompressor:
        p.codec = CODEC["terngrad"]
    elif t is C.NaturalCompressor:
        p.codec = CODEC["natural"]
    elif t is C.U8bitCompressor:
        p.codec = CODEC["u8bit"]
    elif t is C.SignSGDCompressor:
        p.codec = CODEC["signsgd"]
    elif t is C.SignumCompressor:
        p.codec, p.momentum = CODEC["signum"], compressor.momentum
        p.one_minus_momentum = 1.0 - compressor.momentum        # rounded to fp32 from the double, as torch does for a python scalar
    elif t is C.EFSignSGDCompressor:
        p.codec, p.lr = CODEC["efsignsgd"], compressor.learning_rate
    elif t is C.OneBitCompressor:
        p.codec = CODEC["onebit"]
    elif t is C.AdaqCompressor:
        p.codec, p.ratio = CODEC["adaq"], compressor.compress_ratio
    elif t is C.INCEPTIONNCompressor and 0 < compressor.low_exp <= compressor.mid_exp <= 127:
        p.codec, p.quantum_num = CODEC["inceptionn"], int(compressor.low_exp) | (int(compressor.mid_exp) << 8)
    elif t is C.SketchCompressor and 2 <= int(compressor.quantiles) <= 1023:
        p.codec, p.quantum_num = CODEC["sketch"], int(compressor.quantiles)
    else:
        return None
    m = type(memory)
    if m is M.NoneMemory:
        p.memory = MEMORY["none"]
    elif m is M.ResidualMemory:
        p.memory, p.beta, p.gamma = MEMORY["residual"], memory.beta, memory.gamma
    elif m is M.EFSignSGDMemory:
        p.memory, p.beta, p.gamma = MEMORY["residual"], 1.0, memory.learning_rate
    elif m is M.DgcMemory and t is C.DgcCompressor:
        p.memory, p.momentum = MEMORY["dgc"], memory.momentum
        p.clip = 1 if memory.gradient_clipping else 0      # truthy, as in the reference (the dist helper passes 0.3)
    else:
        return None
    return p


def communicator_is_fusable(grace, params):
    """All-gather/broadcast semantics (decode each rank, aggregate) are what the plans
    implement; the all-reduce communicator is equivalent for the linear codecs."""
    name = type(grace).__name__
    mod = type(grace).__module__
    if not mod.startswith("grace_b200."):
        return False
    from grace_b200.ops.native import CODEC
    if name in ("Allgather", "Broadcast"):
        if name == "Broadcast" and not grace.compressor.tensors_size_are_same:
            return False
        return True
    if name == "Allreduce":
        return params.codec in (CODEC["none"], CODEC["fp16"], CODEC["randomk"])
    return False


def _native_or_raise():
    from grace_b200.ops import native
    if not native.available():
        raise native.NativeError(
            "CUDA device present but %s is missing: build it with `bash %s` -- refusing to fall back silently"
            % (native.LIB_PATH, native.BUILD_SH))
    return native


def _grad_view(chunk, p):
    """View of a flat chunk with the parameter's own dense memory format, so autograd's in-place
    accumulation and the optimizer see matching strides (channels_last conv weights)."""
    if p.dim() == 4 and not p.is_contiguous() and p.is_contiguous(memory_format=torch.channels_last):
        return chunk.as_strided(p.shape, p.stride())
    return chunk.view(p.shape)


class _Bucket:
    """Parameters exchanged together through one plan."""

    def __init__(self, engine, named_params, params):
        self.names = [n for n, _ in named_params]
        self.tensors = [p for _, p in named_params]
        self.plan = engine.plan([p.numel() for p in self.tensors], self.names, params)
        self.plans = [self.plan]
        self.flat = torch.zeros(self.plan.padded, dtype=torch.float32, device=self.tensors[0].device)
        self.flats = [self.flat]
        self.views = [_grad_view(self.flat[o:o + p.numel()], p) for o, p in zip(self.plan.offsets, self.tensors)]
        self.pending = len(self.tensors)
        self.sent = False

    def send(self, stream):
        self.plan.send(self.flat, stream=stream)

    def mid(self, stream):
        self.plan.mid(stream=stream)

    def recv(self, stream):
        self.plan.recv(self.flat, stream=stream)

    def receive_all(self, stream):
        """Whole receive side of this bucket on ``stream``: a one-warp flag wait, then the kernels."""
        self.plan.wait(stream=stream)
        self.mid(stream)
        self.recv(stream)


class _PowerSGDBucket(_Bucket):
    """>=2-D gradients go through the native PowerSGD plan (K3), 1-D ones through a dense plan."""

    def __init__(self, engine, named_params, rank, seed, has_memory, use_tc=True, warm=False):
        from grace_b200.ops.native import CodecParams, CODEC, MEMORY
        from grace_b200.parallel.engine import PowerSGDPlan
        self.names = [n for n, _ in named_params]
        self.tensors = [p for _, p in named_params]
        mats = [(n, p) for n, p in named_params if p.dim() >= 2]
        vecs = [(n, p) for n, p in named_params if p.dim() < 2]
        dev = self.tensors[0].device
        self.plans, self.flats = [], []
        view_of = {}
        self.psgd = self.dense = None
        if mats:
            shapes = [(p.shape[0], p.numel() // p.shape[0]) for _, p in mats]
            self.psgd = PowerSGDPlan(engine, shapes, [n for n, _ in mats], rank, seed, has_memory, use_tc, warm)
            self.flat_m = torch.zeros(self.psgd.padded, dtype=torch.float32, device=dev)
            for off, (n, p) in zip(self.psgd.offsets, mats):
                view_of[p] = _grad_view(self.flat_m[off:off + p.numel()], p)
            self.plans.append(self.psgd)
            self.flats.append(self.flat_m)
        if vecs:
            prm = CodecParams()
            prm.codec, prm.memory, prm.average, prm.beta, prm.gamma = CODEC["none"], MEMORY["none"], 1, 1.0, 1.0
            self.dense = engine.plan([p.numel() for _, p in vecs], [n for n, _ in vecs], prm)
            self.flat_v = torch.zeros(self.dense.padded, dtype=torch.float32, device=dev)
            for off, (n, p) in zip(self.dense.offsets, vecs):
                view_of[p] = self.flat_v[off:off + p.numel()].view(p.shape)
            self.plans.append(self.dense)
            self.flats.append(self.flat_v)
        self.views = [view_of[p] for p in self.tensors]
        self.pending = len(self.tensors)
        self.sent = False

    def send(self, stream):
        if self.psgd is not None:
            self.psgd.send(self.flat_m, stream=stream)
        if self.dense is not None:
            self.dense.send(self.flat_v, stream=stream)

    def mid(self, stream):
        if self.psgd is not None:
            self.psgd.mid(stream=stream)
        if self.dense is not None:
            self.dense.mid(stream=stream)

    def recv(self, stream):
        if self.psgd is not None:
            self.psgd.recv(self.flat_m, stream=stream)
        if self.dense is not None:
            self.dense.recv(self.flat_v, stream=stream)

    def receive_all(self, stream):
        if self.psgd is not None:
            self.psgd.wait(0, stream=stream)
            self.psgd.mid(stream=stream)
        if self.dense is not None:
            self.dense.wait(stream=stream)
            self.dense.mid(stream=stream)
        if self.psgd is not None:
            self.psgd.wait(1, stream=stream)
            self.psgd.recv(self.flat_m, stream=stream)
        if self.dense is not None:
            self.dense.recv(self.flat_v, stream=stream)


class _NvlsBucket(_Bucket):
    """Dense (None codec) bucket whose flat gradient buffer is a symmetric allocation bound to an
    NVLS multicast address: the all-reduce is one in-place kernel using multimem.ld_reduce /
    multimem.st (``csrc/k0_dense.cu``).  Construction is collective and may raise when the host does
    not expose multicast; ``FusedDP`` then falls back to the P2P two-shot plan on every rank."""

    def __init__(self, engine, named_params, average):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        from grace_b200.parallel.engine import NvlsPlan
        self.names = [n for n, _ in named_params]
        self.tensors = [p for _, p in named_params]
        offsets, off = [], 0
        for p in self.tensors:
            offsets.append(off)
            off += (p.numel() + 127) // 128 * 128
        dev = self.tensors[0].device
        self.flat = symm.empty(off, dtype=torch.float32, device=dev)
        handle = symm.rendezvous(self.flat, dist.group.WORLD)
        mc = int(handle.multicast_ptr or 0)
        if not mc:
            raise RuntimeError("no NVLS multicast address for this allocation")
        self.handle = handle
        self.flat.zero_()
        self.plan = NvlsPlan(engine, off, mc + int(getattr(handle, "offset", 0) or 0), average)
        self.plans, self.flats = [self.plan], [self.flat]
        self.views = [_grad_view(self.flat[o:o + p.numel()], p) for o, p in zip(offsets, self.tensors)]
        self.pending = len(self.tensors)
        self.sent = False

    def send(self, stream):
        self.plan.signal(stream=stream)          # "my copy is final"; the reduction itself happens in recv

    def mid(self, stream):
        pass

    def recv(self, stream):
        self.plan.reduce(stream=stream)

    def receive_all(self, stream):
        self.plan.reduce(stream=stream)


def _all_ranks_agree(ok):
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()):
        return ok
    t = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return bool(int(t))


def powersgd_spec(grace):
    """(rank, seed, has_memory, warm_start) when ``grace`` is a natively supported PowerSGD triple, else None."""
    from grace_b200.compressor.lowrank import RANK_KEY
    c, m = grace.compressor, grace.memory
    if type(c) is not C.PowerSGDCompressor or type(grace).__name__ != "Allreduce":
        return None
    if not type(grace).__module__.startswith("grace_b200."):
        return None
    if type(m) not in (M.PowerSGDMemory, M.NoneMemory):
        return None
    rank = c.q_memory.get(RANK_KEY, c.rank)
    if not 1 <= rank <= 16:
        return None
    # the previous Q survives to the next step only if the memory does not drop it (memory/powersgd.py:26-28 redraws it)
    warm = bool(c.use_memory) and (type(m) is M.NoneMemory or bool(getattr(m, "warm_start", False)))
    return rank, c.seed, type(m) is M.PowerSGDMemory, warm


class FusedDP:
    """Bucketed, overlapped, native gradient exchange behind ``DistributedOptimizer``."""

    def __init__(self, optimizer, grace, named, params, bucket_bytes=None, powersgd=None):
        self.grace = grace
        import torch.distributed as _dist
        world = _dist.get_world_size() if (_dist.is_available() and _dist.is_initialized()) else 1
        numels = [p.numel() for _, p in named if p.requires_grad]
        self.engine = get_engine(estimate_arena_bytes(numels, params, powersgd, world))
        self.passes = optimizer.backward_passes_per_step
        self.side = torch.cuda.Stream()                   # sends (compensate + compress + peer stores)
        self.side_recv = torch.cuda.Stream()              # receives: queued behind their send, overlap the rest of backward
        self.overlap_recv = os.environ.get("GRACE_B200_OVERLAP_RECV", "1") != "0"
        # "p2p-store": W - 1 peer stores per payload; "nvls-multimem-store": one multicast store through the switch;
        # "nvls-multimem": dense buckets reduced in the switch (multimem.ld_reduce)
        self.transport = "nvls-multimem-store" if getattr(self.engine, "multicast", False) else "p2p-store"
        self.nvls_error = None
        bucket_bytes = bucket_bytes or DEFAULT_BUCKET_BYTES
        trainable = [(n, p) for n, p in named if p.requires_grad]
        # gradients become ready roughly in reverse registration order
        order = list(reversed(trainable))
        # the last bucket's exchange cannot overlap backward (its gradients are the last to arrive): keep it small
        tail_bytes = int(os.environ.get("GRACE_B200_TAIL_BUCKET_MB", "4")) << 20
        total_bytes = sum(p.numel() * 4 for _, p in order)
        tail_start = total_bytes - tail_bytes if total_bytes > 2 * tail_bytes else total_bytes + 1
        self.buckets: List[_Bucket] = []
        cur, cur_bytes, seen = [], 0, 0
        for n, p in order:
            if cur and seen < tail_start <= seen + p.numel() * 4 and cur_bytes >= tail_bytes:
                self.buckets.append(self._make_bucket(cur, params, powersgd))
                cur, cur_bytes = [], 0
            cur.append((n, p))
            cur_bytes += p.numel() * 4
            seen += p.numel() * 4
            if cur_bytes >= bucket_bytes:
                self.buckets.append(self._make_bucket(cur, params, powersgd))
                cur, cur_bytes = [], 0
        if cur:
            self.buckets.append(self._make_bucket(cur, params, powersgd))
        self._where: Dict[torch.Tensor, tuple] = {}
        self._countdown: Dict[torch.Tensor, int] = {}
        self._hooks = []
        # bf16 / fp16 parameters: their .grad keeps the parameter dtype; the exchange (and all error
        # feedback state) runs on an fp32 staging view that is filled when the gradient is final and
        # copied back after the receive
        self._lowp = [p for b in self.buckets for p in b.tensors if p.dtype != torch.float32]
        self._lowp_views = []
        for b in self.buckets:
            for i, p in enumerate(b.tensors):
                if p.dtype != torch.float32:
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
                    self._lowp_views.append(b.views[i])
                else:
                    if p.grad is not None:
                        b.views[i].copy_(p.grad)
                    p.grad = b.views[i]
                self._where[p] = (b, i)
                self._countdown[p] = self.passes
                self._hooks.append(p.register_post_accumulate_grad_hook(self._hook))
        self._inflight = False

    def _make_bucket(self, named_params, params, powersgd):
        from grace_b200.ops.native import CODEC, MEMORY
        if (powersgd is None and self.engine.world > 1 and params.codec == CODEC["none"]
                and params.memory == MEMORY["none"] and os.environ.get("GRACE_B200_NVLS", "1") != "0"
                and not getattr(self, "_nvls_failed", False)):
            bucket, err = None, None
            try:
                bucket = _NvlsBucket(self.engine, named_params, bool(params.average))
            except Exception as exc:                       # no multicast on this host / old driver
                err = exc
            if _all_ranks_agree(bucket is not None):
                self.transport = "nvls-multimem"
                return bucket
            self._nvls_failed = True
            self.nvls_error = repr(err)
        if powersgd is not None:
            rank, seed, has_memory, warm = powersgd
            return _PowerSGDBucket(self.engine, named_params, rank, seed, has_memory,
                                   use_tc=os.environ.get("GRACE_B200_PSGD_TCGEN05", "1") != "0", warm=warm)
        return _Bucket(self.engine, named_params, params)

    # ------------------------------------------------------------------ autograd side
    def _hook(self, p):
        b, i = self._where[p]
        view = b.views[i]
        lowp = p.dtype != torch.float32
        if not lowp and p.grad is not view:
            if p.grad.data_ptr() != view.data_ptr():     # someone replaced .grad (set_to_none): re-attach
                view.copy_(p.grad)
            p.grad = view
        if self._countdown[p] <= 0:
            raise AssertionError("Gradients were computed more than backward_passes_per_step times before call "
                                 "to step(). Increase backward_passes_per_step to accumulate gradients locally.")
        self._countdown[p] -= 1
        self._inflight = True
        if self._countdown[p] == 0:
            if lowp:
                view.copy_(p.grad)                       # final low-precision gradient -> fp32 staging
            b.pending -= 1
            if b.pending == 0:
                self._send(b)

    def _send(self, b):
        ready = torch.cuda.Event()
        ready.record()                                  # on the stream autograd is running on
        self.side.wait_event(ready)
        b.send(self.side)
        b.sent = True
        if self.overlap_recv:
            # receive side right behind the send, on its own stream: a one-warp kernel waits for the peers' flags,
            # then decode + aggregate land in the bucket while backward still runs for the other buckets
            sent = torch.cuda.Event()
            sent.record(self.side)
            self.side_recv.wait_event(sent)
            b.receive_all(self.side_recv)
            b.received = True

    # ------------------------------------------------------------------ optimizer side
    def synchronize(self):
        self.engine.check_status()                      # reports problems of the previous step, no device sync
        for b in self.buckets:
            if not b.sent:
                self._send(b)
        late = [b for b in self.buckets if not getattr(b, "received", False)]
        if late:
            sent = torch.cuda.Event()
            sent.record(self.side)
            self.side_recv.wait_event(sent)
            for b in late:
                b.mid(self.side_recv)                   # PowerSGD / dense two-shot: middle phase of every bucket first
            for b in late:
                b.recv(self.side_recv)                  # aggregate lands in the same views the optimizer reads
        for b in self.buckets:
            b.sent = False
            b.received = False
            b.pending = len(b.tensors)
        done = torch.cuda.Event()
        done.record(self.side_recv)
        torch.cuda.current_stream().wait_event(done)
        if self._lowp:
            torch._foreach_copy_([p.grad for p in self._lowp], self._lowp_views)
        for p in self._countdown:
            self._countdown[p] = self.passes
        self._inflight = False

    def in_flight(self):
        return self._inflight

    def close(self):
        """Detach from the model: remove the hooks, give every parameter a private ``.grad`` again and release the
        plans (their arena slots are not returned -- the arena only grows -- but device state and tables are)."""
        for h in self._hooks:
            h.remove()
        self._hooks = []
        torch.cuda.synchronize()
        for b in self.buckets:
            for p, v in zip(b.tensors, b.views):
                if p.grad is v:
                    p.grad = v.clone()
            for pl in b.plans:
                pl.destroy()
        self.buckets = []

    def zero_grad(self):
        for b in self.buckets:
            for flat in b.flats:
                flat.zero_()
            for p, v in zip(b.tensors, b.views):
                if p.dtype == torch.float32 and p.grad is not v:
                    p.grad = v
        if self._lowp:
            for p in self._lowp:
                if p.grad is None:
                    p.grad = torch.zeros_like(p)
            torch._foreach_zero_([p.grad for p in self._lowp])

    def set_backward_passes_per_step(self, passes):
        self.passes = passes
        for p in self._countdown:
            self._countdown[p] = passes

    def launches(self):
        return sum(pl.launches() for b in self.buckets for pl in b.plans)

    def wire_bytes(self):
        """Bytes this rank stores into each peer per step."""
        return sum(pl.wire_bytes for b in self.buckets for pl in b.plans)

    # ------------------------------------------------------------------ checkpointing
    def state_dict(self):
        torch.cuda.synchronize()
        return {"buckets": [{"names": b.names,
                             "plans": [{"step": pl.step, "state0": pl.get_state(0), "state1": pl.get_state(1)}
                                       for pl in b.plans]} for b in self.buckets]}

    def load_state_dict(self, state):
        """Residual / momentum state is rank-local: save and restore one ``grace_state_dict()`` per rank."""
        torch.cuda.synchronize()
        _restore_buckets(self.buckets, state["buckets"])


def try_build_fused(optimizer, grace, named, required=False, bucket_bytes=None) -> Optional[FusedDP]:
    trainable = [p for _, p in named if p.requires_grad]
    why = None
    if not trainable or not all(p.is_cuda for p in trainable):
        why = "parameters are not on a CUDA device"
    elif not all(p.dtype in (torch.float32, torch.bfloat16, torch.float16) for p in trainable):
        why = "fused path needs float32 / bfloat16 / float16 parameters"
    elif not single_nvlink_domain():
        why = "ranks span several hosts or exceed 8: the peer-memory engine covers one NVLink domain"
    params = None
    psgd = powersgd_spec(grace) if why is None else None
    if why is None and psgd is None:
        params = codec_params(grace.compressor, grace.memory)
        if params is None:
            why = "no native kernels for %s + %s" % (type(grace.compressor).__name__, type(grace.memory).__name__)
        elif not communicator_is_fusable(grace, params):
            why = "%s cannot carry this codec natively" % type(grace).__name__
    if why is not None:
        if required:
            raise RuntimeError("fused=True but " + why)
        return None
    _native_or_raise()
    return FusedDP(optimizer, grace, named, params, bucket_bytes, powersgd=psgd)


class _StepAllSet:
    """Buckets + plans for one fixed list of (name, shape) pairs: the bucketed form of the reference's manual
    ``for name, p in named_parameters(): grc.step(p.grad, name)`` loop (examples/dist/CIFAR10-dawndist/core.py:202-208)."""

    def __init__(self, engine, named_tensors, params, powersgd, bucket_bytes):
        self.engine = engine
        self.buckets = []
        cur, cur_bytes = [], 0
        for n, t in named_tensors:
            cur.append((n, t))
            cur_bytes += t.numel() * 4
            if cur_bytes >= bucket_bytes:
                self.buckets.append(self._make(cur, params, powersgd))
                cur, cur_bytes = [], 0
        if cur:
            self.buckets.append(self._make(cur, params, powersgd))
        for b in self.buckets:
            b.tensors = [None] * len(b.tensors)      # geometry only: do not keep the first call's gradients alive
        self.views = [v for b in self.buckets for v in b.views]

    def _make(self, named_tensors, params, powersgd):
        if powersgd is not None:
            rank, seed, has_memory, warm = powersgd
            return _PowerSGDBucket(self.engine, named_tensors, rank, seed, has_memory,
                                   use_tc=os.environ.get("GRACE_B200_PSGD_TCGEN05", "1") != "0", warm=warm)
        return _Bucket(self.engine, named_tensors, params)

    def exchange(self, grads):
        """Copy the gradients into the buckets (one multi-tensor copy), run every bucket's send, then every mid, then
        every recv on the current stream -- peers' payloads for bucket b arrive while this rank still compresses
        bucket b+1 -- and return the views that now hold the aggregates."""
        torch._foreach_copy_(self.views, grads)
        for b in self.buckets:
            b.send(None)
        for b in self.buckets:
            b.mid(None)
        for b in self.buckets:
            b.recv(None)
        return self.views

    def launches(self):
        return sum(pl.launches() for b in self.buckets for pl in b.plans)

    def destroy(self):
        for b in self.buckets:
            for pl in b.plans:
                pl.destroy()
        self.buckets = []


class FusedStepper:
    """Native fast path of the synchronous ``Communicator.step(tensor, name)`` API and of its bucketed form
    ``Communicator.step_all(named_grads)``."""

    def __init__(self, grace):
        self.grace = grace
        self.params = None
        self.psgd = None
        self.plans = {}
        self.sets = {}
        self.enabled = None
        self._closed_launches = 0

    def _probe(self, tensor):
        if self.enabled is None:
            ok = tensor.is_cuda and os.environ.get("GRACE_B200_FUSED", "1") != "0"
            ok = ok and getattr(self.grace, "group", None) is None      # plans always span the global engine
            if ok:
                self.psgd = powersgd_spec(self.grace)
                if self.psgd is None:
                    self.params = codec_params(self.grace.compressor, self.grace.memory)
                    ok = self.params is not None and communicator_is_fusable(self.grace, self.params)
            ok = ok and single_nvlink_domain()
            if ok:
                _native_or_raise()
                ok = get_engine().world == getattr(self.grace, "world_size", get_engine().world)
            self.enabled = ok
        return self.enabled

    def _entry(self, tensor, name):
        key = (name, tuple(tensor.shape))
        entry = self.plans.get(key)
        if entry is None:
            plan = get_engine().plan([tensor.numel()], [name], self.params)
            entry = (plan, torch.zeros(plan.padded, dtype=torch.float32, device=tensor.device),
                     torch.empty(plan.padded, dtype=torch.float32, device=tensor.device))
            self.plans[key] = entry
            self._apply_state(self._tag(key), plan)
        return entry

    @staticmethod
    def _tag(key):
        return "%s/%s" % (key[0], "x".join(str(d) for d in key[1]))

    def usable(self, tensor):
        # PowerSGD has no single-tensor plan: its native path is step_all (grouped matrices of a bucket)
        return self._probe(tensor) and self.psgd is None and tensor.dtype == torch.float32 and tensor.is_cuda

    def step(self, tensor, name):
        """Returns the aggregated tensor (a fresh tensor, like the reference), or ``None`` when this call must
        take the generic path."""
        if not self.usable(tensor):
            return None
        plan, staging, out = self._entry(tensor, name)
        n = tensor.numel()
        staging[:n].copy_(tensor.reshape(-1))
        plan.send(staging)
        plan.recv(out)
        return out[:n].view(tensor.shape).clone()

    # ---- bucketed manual loop
    def step_all(self, named_grads):
        """``named_grads``: list of (name, tensor).  Returns a list of aggregated tensors (views into persistent
        bucket memory, overwritten by the next call) or ``None`` when the generic path must run."""
        if not named_grads:
            return []
        first = named_grads[0][1]
        if not self._probe(first):
            return None
        if not all(g.is_cuda and g.dtype == torch.float32 for _, g in named_grads):
            return None
        key = tuple((n, tuple(g.shape), tuple(g.stride())) for n, g in named_grads)
        st = self.sets.get(key)
        if st is None:
            st = _StepAllSet(get_engine(), named_grads, self.params, self.psgd, DEFAULT_BUCKET_BYTES)
            self.sets[key] = st
        return st.exchange([g for _, g in named_grads])

    def launches(self):
        return (self._closed_launches + sum(pl.launches() for pl, _, _ in self.plans.values())
                + sum(st.launches() for st in self.sets.values()))

    def close(self):
        torch.cuda.synchronize()
        self._closed_launches = self.launches()
        for plan, _, _ in self.plans.values():
            plan.destroy()
        for st in self.sets.values():
            st.destroy()
        self.plans, self.sets = {}, {}

    def state_dict(self):
        """Error-feedback state of the per-tensor plans and bucket sets (residuals, momenta, step counters)."""
        torch.cuda.synchronize()
        out = {self._tag(key): {"step": plan.step, "state0": plan.get_state(0), "state1": plan.get_state(1)}
               for key, (plan, _, _) in self.plans.items()}
        for i, st in enumerate(self.sets.values()):
            out["__set%d__" % i] = [{"names": b.names, "plans": [{"step": pl.step, "state0": pl.get_state(0),
                                                                  "state1": pl.get_state(1)} for pl in b.plans]}
                                    for b in st.buckets]
        return out

    def load_state_dict(self, state):
        """Restore plans that already exist; entries for tensors not seen yet are kept and applied on
        first use."""
        self._pending_state = dict(state)
        for key, (plan, _, _) in self.plans.items():
            self._apply_state(self._tag(key), plan)
        for i, st in enumerate(self.sets.values()):
            saved = self._pending_state.pop("__set%d__" % i, None)
            if saved is not None:
                _restore_buckets(st.buckets, saved)

    def _apply_state(self, tag, plan):
        st = getattr(self, "_pending_state", {}).pop(tag, None)
        if st is not None:
            plan.step = st["step"]
            for which in (0, 1):
                if st["state%d" % which] is not None:
                    plan.set_state(which, st["state%d" % which])

    # ---- asynchronous halves (Horovod-flavour send_step / receive_step)
    def send(self, tensor, name):
        """Compensate + compress + push on a side stream; returns an opaque ticket or ``None``."""
        if not self.usable(tensor):
            return None
        if not hasattr(self, "_side"):
            self._side = torch.cuda.Stream()
        plan, staging, out = self._entry(tensor, name)
        n = tensor.numel()
        ready = torch.cuda.Event()
        ready.record()
        self._side.wait_event(ready)
        with torch.cuda.stream(self._side):
            staging[:n].copy_(tensor.reshape(-1), non_blocking=True)
        plan.send(staging, stream=self._side)
        return ("native", plan, out, n, tuple(tensor.shape))

    def receive(self, ticket):
        _, plan, out, n, shape = ticket
        plan.recv(out, stream=self._side)
        done = torch.cuda.Event()
        done.record(self._side)
        torch.cuda.current_stream().wait_event(done)
        return out[:n].view(shape).clone()


def _restore_buckets(buckets, saved):
    """Shared by FusedDP and the step_all sets: strict about the layout, so a checkpoint taken with another
    bucket size or model is refused instead of being applied partially."""
    if len(saved) != len(buckets):
        raise ValueError("checkpoint has %d buckets, this run has %d (GRACE_B200_BUCKET_MB changed?)"
                         % (len(saved), len(buckets)))
    for b, s in zip(buckets, saved):
        if s["names"] != b.names:
            raise ValueError("bucket layout differs from the checkpoint")
        if len(s["plans"]) != len(b.plans):
            raise ValueError("bucket %s...: %d plans in the checkpoint, %d here" % (b.names[0], len(s["plans"]), len(b.plans)))
        for pl, ps in zip(b.plans, s["plans"]):
            for which in (0, 1):
                t = ps["state%d" % which]
                if which == 1 and not hasattr(pl, "params"):
                    continue                              # PowerSGD's warm-start Q has its own size; the plan checks it
                if t is not None and t.numel() != pl.padded:
                    raise ValueError("state buffer of bucket %s... has %d elements, the plan expects %d"
                                     % (b.names[0], t.numel(), pl.padded))
    for b, s in zip(buckets, saved):
        for pl, ps in zip(b.plans, s["plans"]):
            pl.step = ps["step"]
            for which in (0, 1):
                if ps["state%d" % which] is not None:
                    pl.set_state(which, ps["state%d" % which])
