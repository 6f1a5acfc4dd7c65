from grace_b200.compressor import OneBitCompressor  # noqa: F401
