from grace_b200.compressor import DgcCompressor  # noqa: F401
