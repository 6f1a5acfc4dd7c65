from grace_b200.compressor import SignumCompressor  # noqa: F401
