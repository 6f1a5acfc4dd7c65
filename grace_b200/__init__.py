"""grace_b200 -- Blackwell-native gradient compression for data-parallel PyTorch.

Same capabilities and public API as GRACE (reference ``grace_dl/__init__.py:1``):
``Communicator`` / ``Compressor`` / ``Memory`` plus ``grace_from_params``.  The hot
path (compensate -> compress -> exchange -> decompress -> aggregate) runs as fused
sm_100a CUDA kernels that store packed payloads straight into peers' memory over
NVLink; the pure-torch classes in this package are the semantic oracle and the
fallback for user-defined codecs.
"""
__version__ = "0.1.0"

from grace_b200.core import Memory, Compressor, Communicator, AsyncCommunicator  # noqa: F401
