from grace_b200.compressor import TernGradCompressor  # noqa: F401
