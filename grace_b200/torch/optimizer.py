"""``DistributedOptimizer`` and the start-of-run broadcast helpers
(reference ``patch_files/horovod/torch/__init__.py:46-250`` and ``:253-403``).

Two execution modes behind one API:

* **fused** (default on CUDA for built-in triples): gradients live in flat buckets that
  are views of engine-owned memory; a bucket's *send* kernel (compensate + compress +
  peer stores) is enqueued on a side stream as soon as autograd has produced the bucket's
  last gradient, so compression and the exchange overlap the rest of backward; the
  *receive* kernels (decode + aggregate, written once into the same views) run in
  ``synchronize()``.  See ``grace_b200/parallel/fused.py``.
* **generic**: one ``grace.send_step`` per parameter from a post-accumulate hook and one
  ``grace.receive_step`` per parameter in ``synchronize()`` -- the reference's behaviour,
  used for user-defined codecs, CPU tensors and Gloo.
"""
from __future__ import annotations

import collections
import warnings
from contextlib import contextmanager

import torch

from grace_b200.torch import mpi_ops as ops


class _DistributedOptimizer(torch.optim.Optimizer):
    def __init__(self, params, named_parameters, grace, backward_passes_per_step=1, fused="auto",
                 bucket_bytes=None):
        super(self.__class__, self).__init__(params)
        self.grace = grace
        every = [p for g in self.param_groups for p in g["params"]]
        if named_parameters is None:
            named = [("allreduce.noname.%d" % i, p) for i, p in enumerate(every)]
        else:
            named = list(named_parameters)
        if any(not isinstance(item, tuple) for item in named):
            raise ValueError("named_parameters should be a sequence of (name, parameter) tuples, "
                             "usually model.named_parameters().")
        counts = collections.Counter(k for k, _ in named)
        dups = sorted(k for k, c in counts.items() if c > 1)
        if dups:
            raise ValueError("Parameter names in named_parameters must be unique. Found duplicates: %s"
                             % ", ".join(dups))
        missing = {id(p) for p in every} - {id(p) for _, p in named}
        if missing:
            raise ValueError("named_parameters was specified, but one or more model parameters were "
                             "not named. Python object ids: %s" % ", ".join(map(str, sorted(missing))))
        self._parameter_names = {p: k for k, p in sorted(named, key=lambda kv: kv[0])}
        self.backward_passes_per_step = backward_passes_per_step
        self._countdown = {p: backward_passes_per_step for p in self._parameter_names}
        self._handles = {}
        self._hook_refs = []
        self._requires_update = set()
        self._synchronized = False
        self._should_synchronize = True
        self._fused = None
        if fused is False and hasattr(grace, "fused"):
            grace.fused = False          # fused=False means: generic path everywhere
        if fused:
            from grace_b200.parallel.fused import try_build_fused
            self._fused = try_build_fused(self, grace, named, required=(fused is True),
                                          bucket_bytes=bucket_bytes)
        if self._fused is None and ops.size() > 1:
            self._register_hooks()
        elif self._fused is None:
            # world size 1: the reference installs no hooks (``__init__.py:89-90``)
            pass

    # ------------------------------------------------------------------ generic path
    def _register_hooks(self):
        for group in self.param_groups:
            for p in group["params"]:
                if p.requires_grad:
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
                    self._requires_update.add(p)
                    self._hook_refs.append(p.register_post_accumulate_grad_hook(self._make_hook()))

    def set_backward_passes_per_step(self, passes):
        self.backward_passes_per_step = passes
        for p in self._countdown:
            self._countdown[p] = passes
        if self._fused is not None:
            self._fused.set_backward_passes_per_step(passes)

    def _send(self, p):
        return self.grace.send_step(p.grad, self._parameter_names.get(p))

    def _make_hook(self):
        def hook(p):
            if p in self._handles and self._handles[p][0] is not None and self._countdown[p] <= 0:
                raise AssertionError(
                    "Gradients were computed more than backward_passes_per_step times before call to "
                    "step(). Increase backward_passes_per_step to accumulate gradients locally.")
            assert self._countdown[p] > 0
            self._countdown[p] -= 1
            self._handles[p] = self._send(p) if self._countdown[p] == 0 else (None, None)
        return hook

    def synchronize(self):
        if self._fused is not None:
            self._fused.synchronize()
            self._synchronized = True
            return
        if ops.size() > 1 or self._handles:
            for p in self._requires_update - set(self._handles):
                self._handles[p] = self._send(p)
            for p, (handles, _) in list(self._handles.items()):
                if handles is None:
                    self._handles[p] = self._send(p)
            for p, (handles, ctx) in self._handles.items():
                out = self.grace.receive_step(handles, ctx)
                self._countdown[p] = self.backward_passes_per_step
                if out.data_ptr() != p.grad.data_ptr():
                    p.grad.copy_(out.view_as(p.grad))
            self._handles.clear()
        self._synchronized = True

    @contextmanager
    def skip_synchronize(self):
        """``optimizer.synchronize(); clip...; with optimizer.skip_synchronize(): optimizer.step()``"""
        self._should_synchronize = False
        try:
            yield
        finally:
            self._should_synchronize = True

    def step(self, closure=None):
        if self._should_synchronize:
            if self._synchronized:
                warnings.warn("optimizer.step() called without optimizer.skip_synchronize() context after "
                              "optimizer.synchronize(). This can cause training slowdown.")
            self.synchronize()
        self._synchronized = False
        if self._fused is not None and not torch.cuda.is_current_stream_capturing():
            # anything the exchange kernels have reported so far (timeout, overflow) stops the update; a plain
            # host read of the pinned status word, no device synchronisation
            self._fused.engine.check_status()
        return super(self.__class__, self).step(closure)

    def zero_grad(self, set_to_none=None):
        if self._handles or (self._fused is not None and self._fused.in_flight()):
            raise AssertionError("optimizer.zero_grad() was called after loss.backward() but before "
                                 "optimizer.step() or optimizer.synchronize(). This is prohibited as it "
                                 "can cause a race condition.")
        if self._fused is not None:
            return self._fused.zero_grad()            # one memset per bucket; views stay attached
        if self._hook_refs:
            return super(self.__class__, self).zero_grad(set_to_none=False)
        return super(self.__class__, self).zero_grad(*(() if set_to_none is None else (set_to_none,)))

    def close(self):
        """Remove the gradient hooks and release the native plans (the optimizer keeps working un-distributed)."""
        for h in self._hook_refs:
            h.remove()
        self._hook_refs = []
        if self._fused is not None:
            self._fused.close()
            self._fused = None

    # ------------------------------------------------------------------ checkpointing
    def grace_state_dict(self):
        """Error-feedback / momentum / step-counter state of the compression pipeline."""
        return self._fused.state_dict() if self._fused is not None else self.grace.state_dict()

    def load_grace_state_dict(self, state):
        (self._fused if self._fused is not None else self.grace).load_state_dict(state)


def DistributedOptimizer(optimizer, grace, named_parameters=None, backward_passes_per_step=1,
                         fused="auto", bucket_bytes=None):
    """Wrap ``optimizer`` so that ``step()`` first exchanges gradients through ``grace``.

    ``grace`` is mandatory, as in the patched Horovod (``horovod_0.18.2.patch:331-352``).
    ``fused``: ``"auto"`` uses the native engine when the (memory, compressor, communicator)
    triple is built-in and the parameters are on CUDA; ``True`` insists; ``False`` forces the
    generic per-parameter path."""
    cls = type(optimizer.__class__.__name__, (optimizer.__class__,), dict(_DistributedOptimizer.__dict__))
    return cls(optimizer.param_groups, named_parameters, grace, backward_passes_per_step, fused, bucket_bytes)


def broadcast_parameters(params, root_rank):
    """In-place broadcast of ``model.state_dict()`` / ``named_parameters()`` / a list."""
    if isinstance(params, dict):
        items = sorted(params.items())
    elif isinstance(params, (list, tuple)) or hasattr(params, "__iter__"):
        items = [p if isinstance(p, tuple) else (None, p) for p in params]
    else:
        raise ValueError("invalid params of type: %s" % type(params))
    handles = [ops.broadcast_async_(p.data if hasattr(p, "data") else p, root_rank, name)
               for name, p in items if torch.is_tensor(p)]
    for h in handles:
        ops.synchronize(h)


def broadcast_optimizer_state(optimizer, root_rank):
    """Make every rank's optimizer state and hyper-parameters equal to ``root_rank``'s."""
    if isinstance(optimizer, torch.optim.LBFGS):
        raise ValueError("cannot broadcast torch.optim.LBFGS state")
    state = optimizer.state_dict()
    if not state["state"]:
        # materialise lazily-created state with a zero-gradient step of the *base* optimizer
        for group in optimizer.param_groups:
            for p in group["params"]:
                if p.requires_grad and p.grad is None:
                    p.grad = torch.zeros_like(p)
        base = optimizer.__class__.__mro__[1] if "_fused" in optimizer.__dict__ else optimizer.__class__
        base.step(optimizer)
        state = optimizer.state_dict()
    if not state["state"]:
        return
    tensors, writers = [], []
    for gi, group in enumerate(state["param_groups"]):
        for key, value in group.items():
            if key == "params" or not isinstance(value, (int, float, bool)):
                continue
            t = torch.tensor([float(value)], dtype=torch.float64)
            tensors.append(("%s.%d" % (key, gi), t))
            # root's value goes into the live group AND into the dict that load_state_dict() applies below
            # (load_state_dict replaces param_groups by the copies taken before the broadcast)
            writers.append(lambda t=t, gi=gi, key=key, kind=type(value):
                           (optimizer.param_groups[gi].__setitem__(key, kind(t.item())),
                            state["param_groups"][gi].__setitem__(key, kind(t.item()))))
        for pid in group["params"]:
            for key, value in sorted(state["state"].get(pid, {}).items(), key=lambda kv: str(kv[0])):
                if torch.is_tensor(value):
                    tensors.append(("%s.%s" % (key, pid), value))
                elif isinstance(value, (int, float, bool)):
                    t = torch.tensor([float(value)], dtype=torch.float64)
                    tensors.append(("%s.%s" % (key, pid), t))
                    writers.append(lambda t=t, pid=pid, key=key, kind=type(value):
                                   state["state"][pid].__setitem__(key, kind(t.item())))
    device = next((v.device for _, v in tensors if v.is_cuda), None)
    moved = [(n, v.to(device) if (device is not None and not v.is_cuda) else v) for n, v in tensors]
    broadcast_parameters(moved, root_rank)
    for (n, orig), (_, sent) in zip(tensors, moved):
        if sent is not orig:
            orig.copy_(sent)
    for w in writers:
        w()
    optimizer.load_state_dict(state)
