from grace_b200.dist.communicator.collectives import Allreduce, Allgather, Broadcast  # noqa: F401
