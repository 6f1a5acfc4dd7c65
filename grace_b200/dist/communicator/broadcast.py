from grace_b200.dist.communicator.collectives import Broadcast  # noqa: F401
