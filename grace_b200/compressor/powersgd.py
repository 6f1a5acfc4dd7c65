from grace_b200.compressor import PowerSGDCompressor  # noqa: F401
