from grace_b200.compressor import RandomKCompressor  # noqa: F401
