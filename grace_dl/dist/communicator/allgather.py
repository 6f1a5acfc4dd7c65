from grace_b200.dist.communicator import Allgather  # noqa: F401
