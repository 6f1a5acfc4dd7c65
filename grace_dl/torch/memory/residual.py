from grace_b200.memory import ResidualMemory  # noqa: F401
