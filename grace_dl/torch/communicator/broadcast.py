from grace_b200.torch.communicator import Broadcast  # noqa: F401
