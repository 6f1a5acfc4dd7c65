"""The three GRACE abstractions (reference ``grace_dl/dist/__init__.py:4-51`` and
``grace_dl/torch/__init__.py:37-58``).

One set of base classes serves both API flavours:

* ``Communicator``       -- synchronous ``step(tensor, name)`` (torch.distributed flavour)
* ``AsyncCommunicator``  -- ``send_step`` / ``receive_step`` (Horovod flavour)

Contract kept from the reference (``IMPLEMENTING.md:3-17``): only the list returned
by ``compress`` crosses the wire, ``ctx`` stays local.  On top of the reference every
object carries ``state_dict`` / ``load_state_dict`` so error-feedback state survives
a checkpoint (checkpoint row).
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Dict

import torch


def _to_plain(v):
    if torch.is_tensor(v):
        return v.detach().clone()
    if isinstance(v, dict):
        return {k: _to_plain(x) for k, x in v.items()}
    return v


class _Stateful:
    """Mixin: snapshot every attribute listed in ``_state_fields``."""

    _state_fields: tuple = ()

    def state_dict(self) -> Dict[str, Any]:
        return {f: _to_plain(getattr(self, f)) for f in self._state_fields}

    def load_state_dict(self, state: Dict[str, Any]) -> None:
        for f in self._state_fields:
            if f in state:
                setattr(self, f, _to_plain(state[f]))


class Memory(_Stateful, ABC):
    """Error-feedback memory: ``compensate`` before compression, ``update`` after."""

    @abstractmethod
    def compensate(self, tensor, name):
        raise NotImplementedError("compensate was not implemented.")

    def update(self, tensor, name, compressor, tensor_compressed, ctx):
        return None


class Compressor(_Stateful, ABC):
    """``average``: divide the aggregate by world size.  ``tensors_size_are_same``:
    payload sizes are rank-invariant, so no size exchange is needed."""

    def __init__(self, average=True, tensors_size_are_same=True):
        self.average = average
        self.tensors_size_are_same = tensors_size_are_same

    @abstractmethod
    def compress(self, tensor, name):
        raise NotImplementedError("compress was not implemented.")

    @abstractmethod
    def decompress(self, tensors, ctx):
        raise NotImplementedError("decompress was not implemented.")

    def aggregate(self, tensors):
        return sum(tensors)


class _CommBase(_Stateful, ABC):
    def state_dict(self):
        out = {"compressor": self.compressor.state_dict(), "memory": self.memory.state_dict()}
        stepper = getattr(self, "_stepper", None)
        if stepper is not None and stepper.plans:           # state held by native per-tensor plans
            out["native"] = stepper.state_dict()
        return out

    def load_state_dict(self, state):
        self.compressor.load_state_dict(state.get("compressor", {}))
        self.memory.load_state_dict(state.get("memory", {}))
        if "native" in state:
            if getattr(self, "_stepper", None) is None:
                from grace_b200.parallel.fused import FusedStepper
                self._stepper = FusedStepper(self)
            self._stepper.load_state_dict(state["native"])

    def _pipeline(self, tensor, name):
        tensor = self.memory.compensate(tensor, name)
        payload, ctx = self.compressor.compress(tensor, name)
        self.memory.update(tensor, name, self.compressor, payload, ctx)
        return payload, ctx


class Communicator(_CommBase):
    """Synchronous flavour (reference ``grace_dl/dist/__init__.py:37-51``)."""

    def __init__(self, compressor, memory, world_size):
        self.compressor = compressor
        self.memory = memory
        self.world_size = world_size

    @abstractmethod
    def send_receive(self, tensors, name, ctx):
        raise NotImplementedError("send_receive was not implemented.")

    def step(self, tensor, name):
        payload, ctx = self._pipeline(tensor, name)
        return self.send_receive(payload, name, ctx)


class AsyncCommunicator(_CommBase):
    """Asynchronous flavour (reference ``grace_dl/torch/__init__.py:37-58``)."""

    def __init__(self, compressor, memory):
        self.compressor = compressor
        self.memory = memory

    @abstractmethod
    def async_send(self, tensors, name):
        raise NotImplementedError("async_send was not implemented.")

    @abstractmethod
    def wait_receive(self, handles, ctx):
        raise NotImplementedError("wait_receive was not implemented.")

    def send_step(self, tensor, name):
        payload, ctx = self._pipeline(tensor, name)
        return self.async_send(payload, name), ctx

    def receive_step(self, handles, ctx):
        return self.wait_receive(handles, ctx)

    # a synchronous convenience so async communicators also drive manual loops
    def step(self, tensor, name):
        handles, ctx = self.send_step(tensor, name)
        return self.receive_step(handles, ctx)
