"""Whole-step CUDA graphs: forward + backward + compressed gradient exchange + optimizer in ONE graph.

At batch 32 a ResNet-50 step on a B200 is partly launch-bound (≈1 500 eager launches for ≈9.5 ms of
GPU work); the reference cannot be captured at all (host syncs in several codecs, one NCCL call per
tensor).  The fused engine can: its kernels read the epoch and the codec step from a device-side
control block advanced by a tick kernel (``csrc/plan.cu``), peers are reached with plain stores and
flag waits inside the kernels, and the side-stream overlap is expressed with events that capture as
graph edges.  Replaying the graph therefore needs no Python, no host-side bookkeeping and no
collective library.

    step = GraphedStep(model, optimizer, loss_fn, example_inputs)   # optimizer: hvd.DistributedOptimizer
    loss = step(x, y)                                               # copies inputs into static buffers, replays

Feeding from the host (data loader with pinned memory) is pipelined the usual way -- the next batch's
host -> device copy runs on a copy stream while the current step executes, and a step's loss comes back
through pinned memory without stalling the launch of the next step:

    step.feed(x0, y0)
    for i in range(n):
        pending = step.run()                    # consumes the fed batch, replays, starts the loss read-back
        if i + 1 < n:
            step.feed(*next_batch())            # overlaps the replay that is in flight
        log(previous.item()) if i else None     # loss of step i-1: already on the host
        previous = pending

(the reference harness copies and synchronises inside the step: examples/torch/pytorch_mnist.py:148-155).
"""
from __future__ import annotations

import torch


class PendingScalar:
    """A device scalar on its way to pinned host memory; ``item()`` waits for the copy only."""

    def __init__(self, host, event, check=None):
        self._host, self._event, self._check = host, event, check

    def done(self):
        return self._event.query()

    def item(self):
        self._event.synchronize()
        if self._check is not None:
            self._check()           # the step that produced this value has finished: report a timeout / overflow now
        return self._host.item()

    def __float__(self):
        return float(self.item())


class PinnedPrefetcher:
    """Double-buffered host -> device input pipeline on a dedicated copy stream.

    ``feed(*host_tensors)`` starts the asynchronous copies of one batch (pinned memory makes them truly
    asynchronous) into one of two device staging sets; ``get()`` makes the current stream wait for the oldest
    fed batch and returns its device tensors; ``release()`` (called by the consumer once it has enqueued
    its reads) lets a later ``feed`` reuse that staging set."""

    def __init__(self, example_inputs, device=None, depth=2):
        device = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.stream = torch.cuda.Stream(device=device)
        self.slots = [[torch.empty(t.shape, dtype=t.dtype, device=device) for t in example_inputs] for _ in range(depth)]
        for slot in self.slots:                      # the staging sets are written on the copy stream: tell the caching
            for t in slot:                            # allocator, so their memory is not handed out again while a copy runs
                t.record_stream(self.stream)
        self.ready = [torch.cuda.Event() for _ in range(depth)]
        self.consumed = [None] * depth
        self.fed, self.taken = 0, 0
        self.bytes_per_batch = sum(t.numel() * t.element_size() for t in example_inputs)

    def feed(self, *host_inputs):
        if self.fed - self.taken >= len(self.slots):
            raise RuntimeError("PinnedPrefetcher: %d batches already in flight" % len(self.slots))
        slot = self.fed % len(self.slots)
        with torch.cuda.stream(self.stream):
            if self.consumed[slot] is not None:
                self.stream.wait_event(self.consumed[slot])       # the previous user of this staging set is done
            for dst, src in zip(self.slots[slot], host_inputs):
                dst.copy_(src, non_blocking=True)
            self.ready[slot].record(self.stream)
        self.fed += 1

    def get(self):
        if self.taken >= self.fed:
            raise RuntimeError("PinnedPrefetcher.get() without a fed batch")
        slot = self.taken % len(self.slots)
        torch.cuda.current_stream().wait_event(self.ready[slot])
        self._current = slot
        self.taken += 1
        return self.slots[slot]

    def release(self):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.consumed[self._current] = ev


def read_back(scalar, host=None, check=None):
    """Start the device -> pinned-host copy of a scalar tensor on the current stream."""
    host = torch.empty((), dtype=scalar.dtype, pin_memory=True) if host is None else host
    host.copy_(scalar, non_blocking=True)
    ev = torch.cuda.Event()
    ev.record(torch.cuda.current_stream())
    return PendingScalar(host, ev, check)


class GraphedStep:
    """``optimizer`` is an ``hvd.DistributedOptimizer`` (its hooks exchange the gradients), or -- for the
    ``torch.distributed`` flavour, whose integration is a manual loop over the parameters
    (examples/dist/CIFAR10-dawndist/core.py:202-208) -- a plain torch optimizer plus ``communicator=``: the captured
    step then runs ``communicator.step_all((name, grad) ...)`` between backward and the optimizer update."""

    def __init__(self, model, optimizer, loss_fn, example_inputs, amp_dtype=torch.bfloat16, warmup=3, communicator=None):
        if not torch.cuda.is_available():
            raise RuntimeError("GraphedStep needs CUDA")
        self.model, self.optimizer, self.loss_fn = model, optimizer, loss_fn
        self.amp_dtype = amp_dtype
        self.communicator = communicator
        self._named = [(n, p) for n, p in model.named_parameters() if p.requires_grad] if communicator is not None else None
        self.static_inputs = [t.clone() for t in example_inputs]
        fused = getattr(optimizer, "_fused", None)
        # eager warm-up on a side stream (cuDNN autotuning, lazy optimizer state, first-call codec rules)
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(max(warmup, 1)):
                self._step_body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        stepper = getattr(communicator, "_stepper", None) if communicator is not None else None
        counter = fused.launches if fused is not None else (stepper.launches if stepper is not None else (lambda: 0))
        n0 = counter()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.static_loss = self._step_body()
        self.native_launches_per_step = counter() - n0
        # Python (and with it FusedDP.synchronize's status check) does not run during a replay: the pinned status
        # word is read around every replay instead -- a plain host load, no device synchronisation
        self._check = None
        if fused is not None:
            self._check = fused.engine.check_status
        elif stepper is not None and stepper.enabled:
            from grace_b200.parallel.fused import get_engine
            self._check = get_engine().check_status
        self.replays = 0
        self._feeder = None
        self._loss_host = None

    def _step_body(self):
        if self.communicator is not None:
            self.optimizer.zero_grad(set_to_none=False)
        else:
            self.optimizer.zero_grad()
        with torch.autocast("cuda", dtype=self.amp_dtype, enabled=self.amp_dtype is not None):
            loss = self.loss_fn(self.model, *self.static_inputs)
        loss.backward()
        if self.communicator is not None:
            self.communicator.step_all([(n, p.grad) for n, p in self._named], inplace=True)
        self.optimizer.step()
        return loss.detach()

    def __call__(self, *inputs):
        for dst, src in zip(self.static_inputs, inputs):
            if src is not dst:
                dst.copy_(src, non_blocking=True)
        if self._check is not None:
            self._check()                                         # problems reported by earlier replays
        self.graph.replay()
        self.replays += 1
        return self.static_loss

    # ---- pipelined host feeding (see the module docstring)
    def feed(self, *host_inputs):
        """Start the host -> device copy of the next step's inputs on the copy stream."""
        if self._feeder is None:
            self._feeder = PinnedPrefetcher(host_inputs, device=self.static_inputs[0].device)
            self._loss_host = [torch.empty((), dtype=self.static_loss.dtype, pin_memory=True) for _ in range(4)]
        self._feeder.feed(*host_inputs)

    def run(self):
        """One step on the oldest fed batch; returns a :class:`PendingScalar` for its loss."""
        staged = self._feeder.get()
        for dst, src in zip(self.static_inputs, staged):
            dst.copy_(src, non_blocking=True)                     # device -> device (+ memory-format change)
        self._feeder.release()
        if self._check is not None:
            self._check()
        self.graph.replay()
        self.replays += 1
        return read_back(self.static_loss, self._loss_host[self.replays % 4], self._check)      # ring of 4: read a loss within 3 steps
