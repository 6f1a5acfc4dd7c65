from grace_b200.compressor.topk import TopKCompressor, sparsify, desparsify  # noqa: F401
