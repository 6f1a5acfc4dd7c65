from grace_b200.compressor import ThresholdCompressor  # noqa: F401
