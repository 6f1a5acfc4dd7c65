from grace_b200.compressor import U8bitCompressor  # noqa: F401
