from grace_b200.memory.memories import EFSignSGDMemory  # noqa: F401
