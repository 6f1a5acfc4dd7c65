"""torch.distributed flavour -- synchronous ``Communicator.step`` API
(reference ``grace_dl/dist/__init__.py``)."""
from grace_b200.core import Memory, Compressor, Communicator  # noqa: F401
from grace_b200.dist.helper import grace_from_params  # noqa: F401


def __getattr__(name):
    # whole-step CUDA graphs and pipelined host feeding for this flavour too (``GraphedStep(..., communicator=grc)``);
    # imported on demand so that ``import grace_b200.dist`` stays free of CUDA-side modules on CPU-only hosts
    if name in ("GraphedStep", "PinnedPrefetcher", "PendingScalar", "read_back"):
        from grace_b200.parallel import graphed
        return getattr(graphed, name)
    raise AttributeError(name)
