from grace_b200.compressor import SketchCompressor  # noqa: F401
