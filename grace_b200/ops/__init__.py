"""Python bindings of ``libgrace_b200.so`` (ctypes over the C ABI in ``include/grace_b200.h``)."""
from grace_b200.ops.native import lib, available, NativeError, check, CodecParams, CODEC, MEMORY  # noqa: F401
