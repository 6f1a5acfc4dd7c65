from grace_b200.torch.communicator import Allreduce  # noqa: F401
