from grace_b200.memory.memories import NoneMemory  # noqa: F401
