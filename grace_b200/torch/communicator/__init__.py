from grace_b200.torch.communicator.collectives import Allreduce, Allgather, Broadcast  # noqa: F401
