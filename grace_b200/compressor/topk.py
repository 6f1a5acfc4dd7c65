from grace_b200.compressor import TopKCompressor  # noqa: F401
