from grace_b200.dist import *  # noqa: F401,F403
from grace_b200.dist import Memory, Compressor, Communicator, grace_from_params  # noqa: F401


def __getattr__(name):      # GraphedStep & co. are resolved lazily by grace_b200.dist (CUDA-side modules)
    import grace_b200.dist as _d
    return getattr(_d, name)
