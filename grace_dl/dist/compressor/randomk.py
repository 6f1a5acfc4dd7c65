from grace_b200.compressor.randomk import RandomKCompressor, sparsify  # noqa: F401
