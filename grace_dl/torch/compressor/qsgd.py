from grace_b200.compressor import QSGDCompressor  # noqa: F401
