"""Workloads built on the runtime - this domain's counterpart of "model families".

The reference ships no workloads beyond its transport tests; BASELINE.json names
two application-level configurations, both implemented here on top of OCM
allocations and our kernels:

* ``gups``  - pooled-HBM random access over a striped allocation (config 3)
* ``sort``  - out-of-core sort of synthetic keys over peer HBM + pinned host (config 4)
"""
from . import gups, sort  # noqa: F401
