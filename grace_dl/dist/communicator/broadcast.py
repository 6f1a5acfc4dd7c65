from grace_b200.dist.communicator import Broadcast  # noqa: F401
