"""Seeded inputs shared by the tests and tests/golden/make_golden.py."""
from cs230_distributed_machine_learning_b200.datasets import make_dataset  # noqa: F401
