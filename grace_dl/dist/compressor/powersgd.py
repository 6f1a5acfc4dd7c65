from grace_b200.compressor.powersgd import PowerSGDCompressor, orthogonalize  # noqa: F401
