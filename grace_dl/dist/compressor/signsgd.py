from grace_b200.compressor import SignSGDCompressor  # noqa: F401
