from grace_b200.compressor import EFSignSGDCompressor  # noqa: F401
