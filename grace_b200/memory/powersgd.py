from grace_b200.memory.memories import PowerSGDMemory  # noqa: F401
