"""Workload models used by examples, tests and benchmarks."""
from grace_b200.models.small import MnistNet, resnet9  # noqa: F401
