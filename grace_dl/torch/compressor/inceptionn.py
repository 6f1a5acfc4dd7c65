from grace_b200.compressor import INCEPTIONNCompressor  # noqa: F401
