from grace_b200.memory.memories import DgcMemory  # noqa: F401
