from grace_b200.dist.helper import grace_from_params  # noqa: F401
