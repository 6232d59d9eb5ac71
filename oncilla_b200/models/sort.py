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
        check: bool = True, gpu_budget_total: int | None = None) -> dict:
    """Sort ``total_keys`` keys using working buffers of ``chunk_keys`` keys.  Returns timings and placement."""
    import torch

    dev = torch.device(f"cuda:{api.device()}")
    n_runs = (total_keys + chunk_keys - 1) // chunk_keys
    n_parts = n_runs
    run_bytes = chunk_keys * 8
    spill = _Spill(gpu_budget_per_rank, gpu_budget_total)
    out_spill = _Spill(gpu_budget_per_rank, gpu_budget_total)
    tiers = {"hbm": 0, "host": 0}
    t_start = time.time()
    # local working set: one OCM window (GPU) big enough for a run / a gathered partition (2x slack)
    work = api.alloc(N.OCM_LOCAL_GPU, 2 * run_bytes)
    tmp = api.alloc(N.OCM_LOCAL_GPU, 2 * run_bytes)
    wt, tt = work.local_tensor(torch.int64), tmp.local_tensor(torch.int64)
    try:
        # splitters from a sample of the synthetic stream (uniform keys -> near-equal partitions)
        sample = torch.cat([_gen(c, 1024, dev) for c in range(n_runs)])
        ss = torch.sort(sample).values
        splitters = ss[torch.arange(1, n_parts, device=dev) * (ss.numel() // n_parts)] if n_parts > 1 else ss[:0]
        runs, bounds, in_sum = [], [], 0
        for c in range(n_runs):
            n = min(chunk_keys, total_keys - c * chunk_keys)
            wt[:n].copy_(_gen(c, n, dev))
            if check:
                in_sum = (in_sum + ops.checksum(wt[:n])) & (2**64 - 1)
            ops.sort_u64(wt[:n], tt[:n])
            b = torch.searchsorted(wt[:n], splitters).tolist()
            bounds.append([0] + b + [n])
            home, tier = spill.place(n * 8)
            tiers[tier] += n * 8
            # K1: local HBM -> far side (peer HBM over NVLink or pinned host)
            api.copy(home, work, n * 8, op_flag=1, stream="sync")
            runs.append((home, n))
        t_runs = time.time()
        outs, out_sum, last_max, sorted_ok = [], 0, -1, True
        for p in range(n_parts):
            fill = 0
            for (home, _n), b in zip(runs, bounds):
                lo, hi = b[p], b[p + 1]
                if hi > lo:
                    if (fill + hi - lo) * 8 > 2 * run_bytes:
                        raise RuntimeError("partition exceeds the working buffer; use more partitions")
                    # K2: slice of a far side -> local window at an offset
                    api.copy(work, home, (hi - lo) * 8, src_offset=lo * 8, dest_offset=fill * 8, op_flag=1)
                    fill += hi - lo
            if fill == 0:
                outs.append((None, 0))
                continue
            ops.sort_u64(wt[:fill], tt[:fill])
            if check:
                part = wt[:fill]
                sorted_ok &= bool((part[1:] >= part[:-1]).all().item()) and int(part[0].item()) >= last_max
                last_max = int(part[-1].item())
                out_sum = (out_sum + ops.checksum(part)) & (2**64 - 1)
            dst, tier = out_spill.place(fill * 8)
            api.copy(dst, work, fill * 8, op_flag=1)
            outs.append((dst, fill))
        torch.cuda.synchronize()
        t_end = time.time()
        total_out = sum(n for _, n in outs)
        return {"total_keys": total_keys, "n_runs": n_runs, "seconds": t_end - t_start,
                "phase1_s": t_runs - t_start, "phase2_s": t_end - t_runs,
                "mkeys_per_s": total_keys / (t_end - t_start) / 1e6, "bytes_in_hbm": tiers["hbm"],
                "bytes_in_host": tiers["host"],
                "ok": (not check) or (sorted_ok and out_sum == in_sum and total_out == total_keys)}
    finally:
        work.free()
        tmp.free()
        spill.free()
        out_spill.free()
