from grace_b200.torch.communicator import Allgather  # noqa: F401
