from grace_b200.memory import EFSignSGDMemory  # noqa: F401
