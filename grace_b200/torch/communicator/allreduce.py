from grace_b200.torch.communicator.collectives import Allreduce  # noqa: F401
