from grace_b200.dist.communicator.collectives import Allgather  # noqa: F401
