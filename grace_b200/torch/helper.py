"""``grace_from_params`` for the Horovod-style flavour (reference
``grace_dl/torch/helper.py:1-80``).  The reference hard-codes every hyper-parameter there;
the same values are the *defaults* here and any key present in ``params`` overrides them."""
from __future__ import annotations

from grace_b200.registry import build_compressor, build_memory


def grace_from_params(params):
    from grace_b200.torch import mpi_ops as ops
    from grace_b200.torch.communicator import Allreduce, Allgather, Broadcast

    world_size = params.get("world_size", ops.size())
    compressor = build_compressor(params, "torch", world_size)
    if hasattr(compressor, "_allreduce") and type(compressor).__name__ == "PowerSGDCompressor":
        compressor._allreduce = lambda t: ops.allreduce_(t, average=True)
    memory = build_memory(params, "torch", world_size, compressor)
    if type(memory).__name__ == "DgcMemory":
        memory._allreduce_sum = lambda t: ops.allreduce_(t, average=False)
    kind = params.get("communicator", "allreduce")
    if kind == "allreduce":
        return Allreduce(compressor, memory)
    if kind == "allgather":
        return Allgather(compressor, memory, world_size)
    if kind == "broadcast":
        return Broadcast(compressor, memory, world_size)
    raise NotImplementedError(kind)
