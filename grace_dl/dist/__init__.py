from grace_b200.dist import *  # noqa: F401,F403
from grace_b200.dist import Memory, Compressor, Communicator, grace_from_params  # noqa: F401
