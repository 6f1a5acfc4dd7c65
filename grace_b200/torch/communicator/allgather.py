from grace_b200.torch.communicator.collectives import Allgather  # noqa: F401
