"""torch.distributed flavour -- synchronous ``Communicator.step`` API
(reference ``grace_dl/dist/__init__.py``)."""
from grace_b200.core import Memory, Compressor, Communicator  # noqa: F401
from grace_b200.dist.helper import grace_from_params  # noqa: F401
