"""Horovod-style flavour: ``import grace_b200.torch as hvd`` gives the API the reference's
patched Horovod exposes (``patch_files/horovod/torch/__init__.py:32-41,204-403``) plus the
asynchronous GRACE abstractions (``grace_dl/torch/__init__.py:4-58``)."""
from grace_b200.core import Memory, Compressor  # noqa: F401
from grace_b200.core import AsyncCommunicator as Communicator  # noqa: F401
from grace_b200.torch.mpi_ops import (  # noqa: F401
    init, shutdown, size, rank, local_rank, local_size, poll, synchronize,
    allreduce, allreduce_async, allreduce_, allreduce_async_,
    allgather, allgather_async, broadcast, broadcast_async, broadcast_, broadcast_async_,
    mpi_threads_supported, nccl_built, gloo_built)
from grace_b200.torch.optimizer import (  # noqa: F401
    DistributedOptimizer, broadcast_parameters, broadcast_optimizer_state)
from grace_b200.torch.helper import grace_from_params  # noqa: F401
from grace_b200.parallel.graphed import GraphedStep, PinnedPrefetcher, PendingScalar, read_back  # noqa: F401,E402
