from grace_b200.torch import *  # noqa: F401,F403
from grace_b200.torch import Memory, Compressor, Communicator, grace_from_params  # noqa: F401
