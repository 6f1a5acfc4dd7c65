from grace_b200.torch.communicator.collectives import Broadcast  # noqa: F401
