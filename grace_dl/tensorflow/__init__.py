"""``grace_dl.tensorflow`` -- not provided.

The reference's third flavour (``grace_dl/tensorflow/__init__.py:6-108`` and the Horovod-TF patch) builds TensorFlow
graph ops.  TensorFlow is outside this framework's scope (BASELINE.json) and is not installed in the target image,
so importing this package raises instead of failing later with an obscure error.  Everything the TF flavour had that
the torch flavours lacked -- the AdaQ, SketchML, 8-bit and INCEPTIONN codecs and their helper defaults -- is available
from ``grace_b200.torch`` / ``grace_b200.dist`` (and the ``grace_dl.torch`` / ``grace_dl.dist`` shims), with native
sm_100a kernels.
"""
raise ImportError(
    "grace_dl.tensorflow is not part of grace_b200 (PyTorch-only, TensorFlow is out of scope). Use grace_dl.torch "
    "(Horovod-style API) or grace_dl.dist (torch.distributed-style API); the TF-only codecs 'adaq', 'sketch', 'u8bit' "
    "and 'inceptionn' are accepted by both grace_from_params helpers.")
