from grace_b200.memory.memories import ResidualMemory  # noqa: F401
