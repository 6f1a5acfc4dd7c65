from grace_b200.memory import DgcMemory  # noqa: F401
