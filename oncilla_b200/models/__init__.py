"""Workloads built on the runtime - this domain's counterpart of "model families".

The reference ships no workloads beyond its transport tests; BASELINE.json names
two application-level configurations, both implemented here on top of OCM
allocations and our kernels:

* ``gups``  - pooled-HBM random access over a striped allocation (config 3)
* ``sort``  - out-of-core sort of synthetic keys over peer HBM + pinned host (config 4)
* ``dsort`` - distributed sample sort: every rank sorts, keys are exchanged with one-sided puts into
              blocks shared between the applications (``ocm_share`` / ``ocm_attach``) and per-block signals
"""
from . import dsort, gups, sort  # noqa: F401
