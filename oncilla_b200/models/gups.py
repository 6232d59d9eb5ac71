"""GUPS-style random access over pooled HBM (BASELINE.json config 3).

One process drives its GPU; the table is a ``StripedTable`` whose slabs live in
every rank's HBM pool.  Updates are 8-byte xor reductions issued from inside the
kernel directly to the owning GPU over NVLink (kernel K7, src/kernels/gups.cu).
Applying the same update stream twice restores the table, which is the check.
"""
from __future__ import annotations

from .. import ops
from ..parallel import StripedTable


def run(total_words: int = 1 << 26, n_updates: int = 1 << 26, words_per_stripe: int = 1 << 16, seed: int = 7,
        iters: int = 3, verify: bool = True, local_only: bool = False) -> dict:
    import torch

    from .. import api

    # local_only: the whole table in this rank's own HBM (no NVLink) - the reference point for the pooled number
    tab = StripedTable(total_words, words_per_stripe, ranks=[api.rank()] if local_only else None)
    try:
        # far sides start zeroed; keep it that way (xor stream applied twice must give zeros back)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ops.gups(tab.bases, tab.words_per_stripe, total_words, n_updates, seed)  # warm-up, applied once
        torch.cuda.synchronize()
        times = []
        for _ in range(iters):
            ev0.record()
            ops.gups(tab.bases, tab.words_per_stripe, total_words, n_updates, seed)
            ev1.record()
            torch.cuda.synchronize()
            times.append(ev0.elapsed_time(ev1) / 1e3)
        applied = 1 + iters
        ok = None
        if verify:
            if applied % 2:
                ops.gups(tab.bases, tab.words_per_stripe, total_words, n_updates, seed)
                torch.cuda.synchronize()
            ok = all(ops.verify(tab.slab_tensor(r), 0) == 0 for r in range(tab.n_ranks))
        best = min(times)
        return {"gups": n_updates / best / 1e9, "seconds": best, "total_words": total_words, "n_updates": n_updates,
                "n_ranks": tab.n_ranks, "verified": ok, "my_rank": api.rank()}
    finally:
        tab.free()
