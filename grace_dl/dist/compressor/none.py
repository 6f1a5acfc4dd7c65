from grace_b200.compressor import NoneCompressor  # noqa: F401
