from grace_b200.dist.communicator.collectives import Allreduce  # noqa: F401
