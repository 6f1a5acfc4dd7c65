from grace_b200.memory import NoneMemory  # noqa: F401
