from grace_b200.dist.communicator import Allreduce  # noqa: F401
