"""Helpers for the benchmark harness: CPU partitioning and machine-state sampling."""
from .cpus import available_cpus, split_cpus  # noqa: F401
from .clocks import ClockSampler, cpu_info  # noqa: F401
