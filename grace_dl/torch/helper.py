from grace_b200.torch.helper import grace_from_params  # noqa: F401
