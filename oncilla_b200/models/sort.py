"""Out-of-core sort of synthetic 64-bit keys (BASELINE.json config 4).

The data set does not fit the driving GPU's working buffers: runs and the
sorted output live in far-side OCM allocations spread round-robin over every
rank's HBM pool and, once those are full, over the pinned host-DRAM tier.

  phase 1  run formation: generate a chunk, radix-sort it on the GPU (K8),
           record where each splitter falls, put the run to its home (K1)
  phase 2  per partition: get the matching slice of every run (K2, offsets into
           the far sides), radix-sort the gathered partition, put it to the output
  check    every partition is sorted, partition boundaries are ordered and the
           wrapping key checksum (K9) equals the input's

Keys are < 2**63 so that torch's signed comparisons agree with the unsigned sort.
"""
from __future__ import annotations

import time

from .. import _native as N
from .. import api, ops
from ..parallel import alloc_on


class _Spill:
    """Far-side storage that prefers peer HBM and falls back to the pinned host tier."""

    def __init__(self, gpu_budget_per_rank: int, gpu_budget_total: int | None = None):
        self.n_ranks = api.num_ranks()
        self.left = [gpu_budget_per_rank] * self.n_ranks
        self.total_left = gpu_budget_total if gpu_budget_total is not None else 1 << 62
        self.next = 0
        self.allocs = []

    def place(self, nbytes: int):
        for _ in range(self.n_ranks):
            r = self.next
            self.next = (self.next + 1) % self.n_ranks
            if self.left[r] >= nbytes and self.total_left >= nbytes:
                try:
                    a = alloc_on(r, nbytes, gpu=True, flags=N.ALLOC_NO_ZERO)
                    if a.kind == N.OCM_REMOTE_GPU:
                        self.left[r] -= nbytes
                        self.total_left -= nbytes
                        self.allocs.append(a)
                        return a, "hbm"
                    a.free()
                except N.OcmError:
                    pass
        r = self.next
        self.next = (self.next + 1) % self.n_ranks
        a = alloc_on(r, nbytes, gpu=False, flags=N.ALLOC_NO_ZERO)
        self.allocs.append(a)
        return a, "host"

    def free(self):
        for a in self.allocs:
            a.free()
        self.allocs = []


def _gen(chunk_index: int, n: int, device):
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(0x5EED + chunk_index)
    return torch.randint(0, 2**62, (n,), dtype=torch.int64, device=device, generator=g)


def run(total_keys: int = 1 << 24, chunk_keys: int = 1 << 21, gpu_budget_per_rank: int = 1 << 62,
        check: bool = True, gpu_budget_total: int | None = None, final: str = "merge") -> dict:
    """One process sorts ``total_keys`` keys alone with working buffers of ``chunk_keys`` keys: the single-rank form of
    the pipeline in :mod:`oncilla_b200.models.oocsort` (two streams, runs put while the next one is sorted, partitions
    merged from sorted slices).  ``gpu_budget_total`` bytes of runs + output may live in peer HBM, the rest spills to
    the pinned host tier."""
    from . import oocsort

    budget = gpu_budget_total * 2 if gpu_budget_total is not None else gpu_budget_per_rank
    r = oocsort.run(keys_per_rank=total_keys, chunk_keys=chunk_keys, hbm_budget=budget, final=final, check=check,
                    collaborate=False)
    r["total_keys"] = total_keys
    return r
