"""``grace_from_params`` for the torch.distributed flavour
(reference ``grace_dl/dist/helper.py:1-86``: same keys, same defaults).

Extras: ``'efsignsgd'`` memory and a working ``'broadcast'``, the TF-only
codecs (``adaq``, ``sketch``, ``u8bit``/``8bit``, ``inceptionn``), and ``params['fused']``
(default: auto) which selects the native sm_100a engine for CUDA tensors.
"""
from __future__ import annotations

from torch import distributed as dist

from grace_b200.registry import build_compressor, build_memory


def grace_from_params(params):
    from grace_b200.dist.communicator import Allreduce, Allgather, Broadcast

    if "world_size" in params:
        world_size = params["world_size"]
    elif dist.is_available() and dist.is_initialized():
        world_size = dist.get_world_size()
    else:
        raise KeyError("world_size")
    compressor = build_compressor(params, "dist", world_size)
    memory = build_memory(params, "dist", world_size, compressor)
    kind = params.get("communicator", "allreduce")
    if kind == "allreduce":
        return Allreduce(compressor, memory, world_size)
    if kind == "allgather":
        return Allgather(compressor, memory, world_size)
    if kind == "broadcast":
        return Broadcast(compressor, memory, world_size, rank=params.get("rank"))
    raise NotImplementedError(kind)
