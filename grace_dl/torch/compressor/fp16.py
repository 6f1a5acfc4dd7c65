from grace_b200.compressor import FP16Compressor  # noqa: F401
