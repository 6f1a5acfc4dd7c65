from grace_b200.compressor import NaturalCompressor  # noqa: F401
