"""Placement and striping: the "parallelism strategies" of a memory runtime.

The reference has exactly one policy - the neighbour rank ``(r+1) % N`` (reference
src/alloc.c:107,120) - and lists striping one allocation over several nodes as a
TODO (reference src/lib.c:42-45).  Here policies are selectable per allocation and
``StripedTable`` spreads one logical array over every rank's HBM pool.
"""
from __future__ import annotations

from .. import _native as N
from ..api import Allocation

POLICIES = {
    "default": N.POLICY_DEFAULT,
    "neighbour": N.POLICY_NEIGHBOUR,
    "explicit": N.POLICY_EXPLICIT,
    "least_loaded": N.POLICY_LEAST_LOADED,
    "round_robin": N.POLICY_ROUND_ROBIN,
    "self": N.POLICY_SELF,
}


def alloc_on(rank: int, nbytes: int, *, gpu: bool = True, window_bytes: int = 0, flags: int = 0) -> Allocation:
    """A far-side region on an explicit rank (HBM if ``gpu`` else pinned host)."""
    kind = N.OCM_REMOTE_GPU if gpu else N.OCM_REMOTE_HOST
    if not window_bytes:
        flags |= N.ALLOC_NO_WINDOW
    return Allocation(kind, window_bytes, nbytes, policy=N.POLICY_EXPLICIT, rank=rank, flags=flags)


class StripedTable:
    """One logical array of 8-byte words striped over the HBM pools of all ranks.

    Word ``i`` lives on rank ``(i // words_per_stripe) % n_ranks``.  Every slab is mapped
    into the calling process, so device code addresses any word directly (K7 GUPS).
    """

    def __init__(self, total_words: int, words_per_stripe: int = 1 << 16, n_ranks: int | None = None,
                 ranks: list[int] | None = None):
        from .. import api

        assert total_words & (total_words - 1) == 0 and words_per_stripe & (words_per_stripe - 1) == 0
        # ranks: the ranks that hold slab 0, 1, ... (default: all of them in order; [api.rank()] = one local HBM)
        self.ranks = list(ranks) if ranks is not None else list(range(n_ranks or api.num_ranks()))
        self.n_ranks = len(self.ranks)
        self.total_words = total_words
        self.words_per_stripe = min(words_per_stripe, total_words)
        stripes = total_words // self.words_per_stripe
        per_rank_stripes = (stripes + self.n_ranks - 1) // self.n_ranks
        self.slab_bytes = per_rank_stripes * self.words_per_stripe * 8
        self.slabs = [alloc_on(r, self.slab_bytes) for r in self.ranks]

    @property
    def bases(self) -> list[int]:
        return [s.remote_ptr for s in self.slabs]

    def slab_tensor(self, r: int):
        import torch

        return self.slabs[r].remote_tensor(torch.int64)

    def locate(self, idx: int) -> tuple[int, int]:
        stripe, within = divmod(idx, self.words_per_stripe)
        return stripe % self.n_ranks, (stripe // self.n_ranks) * self.words_per_stripe + within

    def free(self) -> None:
        for s in self.slabs:
            s.free()
        self.slabs = []
