from grace_b200.compressor import AdaqCompressor  # noqa: F401
