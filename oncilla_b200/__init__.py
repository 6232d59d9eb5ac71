"""oncilla_b200 - Python binding and harness of the OCM-B200 runtime.

The runtime itself is native (C++/CUDA for sm_100a, built under ``build/``):
``oncillamem`` daemons own per-rank HBM and pinned-host pools, ``libocm.so``
implements the ``ocm_*`` API and the data-movement kernels.  This package only
binds them:

* ``api``      - ``init/tini``, ``Allocation``, ``copy``, ``copy_strided``, ``stats``
* ``cluster``  - start/stop daemons on this host
* ``ops``      - the kernels on raw torch tensors, plus the CPU oracles
* ``parallel`` - placement policies and striped allocations
* ``models``   - workloads: GUPS over pooled HBM, out-of-core sort
* ``utils``    - clocks sampling, L2 flush
"""
from . import _native
from ._native import (ALLOC_NO_WINDOW, ALLOC_NO_ZERO, ALLOC_SLOW_PATH, DT_BF16, DT_F32, DT_F8E4M3, DT_U8,
                      OCM_LOCAL_GPU, OCM_LOCAL_HOST, OCM_REMOTE_GPU, OCM_REMOTE_HOST, OCM_REMOTE_RDMA, OCM_REMOTE_RMA,
                      OCM_REPLICATED_GPU, OCM_STRIPED_GPU, POLICY_DEFAULT, POLICY_EXPLICIT, POLICY_LEAST_LOADED,
                      POLICY_NEIGHBOUR, POLICY_ROUND_ROBIN, POLICY_SELF, OcmError, build_native)
from .api import (Allocation, alloc, attach, copy, copy_strided, device, init, multicast_ready, num_ranks,
                  prefetch_views, quiet, rank, server_stats, stats, tini)
from .cluster import LocalCluster, write_nodefile

__all__ = [n for n in dir() if not n.startswith("_")]
