from grace_b200.memory import PowerSGDMemory  # noqa: F401
