"""Out-of-core sort of synthetic 64-bit keys as a pipeline over pooled HBM + the pinned host tier (BASELINE.json
config 4).  Every rank runs it at once; ``torch.distributed`` carries only the control plane (samples,
slice bounds, 32-byte share tokens), every key moves through OCM allocations and our kernels.

  phase 1  run formation, double-buffered on two streams: while run c is being PUT to its home (K1, a far side in
           some rank's HBM pool or - once the HBM budget is spent - in a pinned host pool), run c+1 is generated and
           radix-sorted (K8) in the other working buffer.  Where the global splitters fall in each run is recorded
           on the device and read once, after the loop.
  control  every rank shares its runs (``ocm_share``), all ranks exchange tokens and slice bounds.
  phase 2  per partition (each rank owns ``parts_per_rank`` consecutive partitions of the global order), again on two
           streams: GET the partition's slice of EVERY run of EVERY rank (K2, through ``ocm_attach`` handles) into the
           working buffer - the slices are sorted, so k-way MERGING them (merge-path kernel, log2 k levels) finishes
           the partition without re-sorting - and PUT the result to its home.  The gets of partition q+1 overlap the
           merge of partition q.
  check    every partition ascending, partitions ordered across ranks, wrapping key checksum equal to the input's.

Keys use the full unsigned 64-bit range.  Nothing here calls NCCL or a library copy for key data.
"""
from __future__ import annotations

import time

from .. import _native as N
from .. import api, ops
from ..parallel import alloc_on


def _dist():
    import torch.distributed as dist

    return dist if dist.is_available() and dist.is_initialized() else None


class Spill:
    """Homes for runs / output: peer HBM round-robin over all ranks while the budget lasts, then pinned host pools."""

    def __init__(self, hbm_budget: int, start_rank: int = 0):
        self.n_ranks = api.num_ranks()
        self.hbm_left = int(hbm_budget)
        self.next = start_rank % self.n_ranks
        self.allocs: list = []
        self.bytes = {"hbm": 0, "host": 0}

    def place(self, nbytes: int):
        for want_gpu in (True, False):
            if want_gpu and self.hbm_left < nbytes:
                continue
            for _ in range(self.n_ranks):
                r = self.next
                self.next = (self.next + 1) % self.n_ranks
                try:
                    a = alloc_on(r, nbytes, gpu=want_gpu, flags=N.ALLOC_NO_ZERO)
                except N.OcmError:
                    continue
                if want_gpu and a.kind != N.OCM_REMOTE_GPU:      # the governor down-tiered it: not what we budgeted
                    a.free()
                    continue
                tier = "hbm" if want_gpu else "host"
                if want_gpu:
                    self.hbm_left -= nbytes
                self.bytes[tier] += nbytes
                self.allocs.append(a)
                return a, tier
        raise N.OcmError(f"no room for {nbytes} bytes in any tier")

    def free(self):
        for a in self.allocs:
            a.free()
        self.allocs = []


def _gen_into(t, rank: int, chunk_index: int, device):
    """Synthetic keys for (rank, chunk): full-range 64-bit values, reproducible."""
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(0x5EED + rank * 1_000_003 + chunk_index)
    t.random_(-2**63, 2**63 - 1, generator=g)
    return t


def _u(t):
    """Map unsigned key order onto torch's signed comparisons (flip the sign bit)."""
    import torch

    return t ^ torch.tensor(-2**63, dtype=torch.int64, device=t.device)


def run(keys_per_rank: int = 1 << 22, chunk_keys: int = 1 << 20, parts_per_rank: int | None = None,
        hbm_budget: int = 1 << 62, final: str = "merge", check: bool = True, collaborate: bool = True,
        prefetch: bool = True) -> dict:
    """Sort ``keys_per_rank`` synthetic keys per rank into one global order (rank r ends up owning partitions
    ``r * parts_per_rank ...``).  ``chunk_keys`` keys fit one working buffer; ``hbm_budget`` bytes of this rank's runs
    and output may live in peer HBM, the rest spills to pinned host pools.  ``final``: "merge" (k-way merge of the
    gathered sorted slices) or "sort" (radix sort of the gathered partition, the round-1 behaviour, for comparison).
    ``collaborate=False`` ignores ``torch.distributed``: this process sorts its own keys alone.
    ``prefetch``: map every rank's pools before the clock starts (once per process; reported as ``setup_s``) - the
    sort touches all of them, and mapping seven 100 GiB peer pools on first use cost 29 s of a 41 s first pass."""
    import torch

    dist = _dist() if collaborate else None
    rank = dist.get_rank() if dist else 0
    world = dist.get_world_size() if dist else 1
    dev = torch.device(f"cuda:{api.device()}")
    n_runs = (keys_per_rank + chunk_keys - 1) // chunk_keys
    S = parts_per_rank or n_runs                     # partitions per rank: about one working buffer each
    P = world * S
    buf_keys = 2 * chunk_keys                        # working buffers hold a run, or a gathered partition with slack
    streams = [torch.cuda.Stream(device=dev) for _ in range(2)]
    work = [api.alloc(N.OCM_LOCAL_GPU, buf_keys * 8, flags=N.ALLOC_NO_ZERO) for _ in range(2)]
    tmp = [api.alloc(N.OCM_LOCAL_GPU, buf_keys * 8, flags=N.ALLOC_NO_ZERO) for _ in range(2)]
    wt = [w.local_tensor(torch.int64) for w in work]
    tt = [t.local_tensor(torch.int64) for t in tmp]
    runs_spill, out_spill = Spill(hbm_budget // 2, rank), Spill(hbm_budget // 2, rank + 1)
    attached: list = []
    host_s = {"place": 0.0, "copy_calls": 0.0, "control": 0.0, "splitters": 0.0}   # host time spent blocked in each kind of call

    def timed(key, fn, *a, **kw):
        t = time.perf_counter()
        r = fn(*a, **kw)
        host_s[key] += time.perf_counter() - t
        return r

    def all_gather_obj(obj):
        if not dist:
            return [obj]
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()

    def add_checksum(acc, t):
        # wrapping sum of the keys, accumulated on the device (no host round trip inside the pipeline)
        ops._check(ops.lib.ocmk_checksum(ops.C.c_void_p(t.data_ptr()), t.numel() * 8, ops.C.c_void_p(acc.data_ptr()),
                                         ops._stream()), "ocmk_checksum")

    t_setup = time.time()
    if prefetch:
        api.prefetch_views()
        if dist:
            dist.barrier()
    setup_s = time.time() - t_setup
    try:
        # ---- splitters: a sample of every rank's stream -> P - 1 global splitters (the same on every rank) ----------
        t_start = time.time()
        mine = torch.cat([_gen_into(torch.empty(1024, dtype=torch.int64, device=dev), rank, 1_000_000 + c, dev)
                          for c in range(min(n_runs, 64))])
        samples = torch.cat([s.to(dev) for s in all_gather_obj(mine.cpu())])
        ss = torch.sort(_u(samples)).values                       # signed view of the unsigned order
        spl = ss[torch.arange(1, P, device=dev) * (ss.numel() // P)] if P > 1 else ss[:0]

        host_s["splitters"] = time.time() - t_start
        # ---- phase 1: run formation, two buffers / two streams -------------------------------------------------
        in_acc = torch.zeros(1, dtype=torch.int64, device=dev)
        bounds_dev = torch.zeros(n_runs, P + 1, dtype=torch.int64, device=dev)
        runs, sort_done = [], None
        for c in range(n_runs):
            b = c & 1
            n = min(chunk_keys, keys_per_rank - c * chunk_keys)
            home, _tier = timed("place", runs_spill.place, n * 8)
            with torch.cuda.stream(streams[b]):
                keys = wt[b][:n]
                _gen_into(keys, rank, c, dev)
                if check:
                    add_checksum(in_acc, keys)
                if sort_done is not None:
                    streams[b].wait_event(sort_done)              # the sort's scratch is one per GPU: sorts take turns
                ops.sort_u64(keys, tt[b][:n])
                sort_done = torch.cuda.Event()
                sort_done.record(streams[b])
                bounds_dev[c, 1:P] = torch.searchsorted(_u(keys), spl)
                bounds_dev[c, P] = n
                # K1: the run leaves for its home while the other buffer is being generated and sorted
                timed("copy_calls", api.copy, home, work[b], n * 8, stream=streams[b])
            runs.append((home, n))
        barrier()
        t_runs = time.time()

        # ---- control plane: who holds which run, and where the splitters cut it ------------------------------------
        t_ctl = time.perf_counter()
        my_bounds = bounds_dev.cpu().tolist()
        info = all_gather_obj({"tokens": [h.share().hex() for h, _ in runs], "bounds": my_bounds})
        sources = []                                              # (handle, bounds) of every run of every rank
        for r, inf in enumerate(info):
            for i, (tok, bnd) in enumerate(zip(inf["tokens"], inf["bounds"])):
                if r == rank:
                    h = runs[i][0]
                else:
                    h = api.attach(bytes.fromhex(tok))
                    attached.append(h)
                sources.append((h, bnd))

        host_s["control"] = time.perf_counter() - t_ctl
        # ---- phase 2: my partitions; gets of partition q+1 overlap the merge of partition q ------------------------
        out_acc = torch.zeros(1, dtype=torch.int64, device=dev)
        ok_dev = torch.ones(1, dtype=torch.bool, device=dev)
        edges = torch.zeros(S, 2, dtype=torch.int64, device=dev)   # first / last key of each partition (signed view)
        sizes, outs, merge_done = [], [], None
        for q in range(S):
            p = rank * S + q
            b = q & 1
            offs, fill = [0], 0
            with torch.cuda.stream(streams[b]):
                for h, bnd in sources:
                    lo, hi = bnd[p], bnd[p + 1]
                    if hi > lo:
                        if fill + hi - lo > buf_keys:
                            raise RuntimeError("partition exceeds the working buffer: use more partitions per rank")
                        # K2: this partition's slice of a run, from wherever the run lives, to its place in the buffer
                        timed("copy_calls", api.copy, work[b], h, (hi - lo) * 8, src_offset=lo * 8, dest_offset=fill * 8,
                              stream=streams[b])
                        fill += hi - lo
                        offs.append(fill)
                sizes.append(fill)
                if fill == 0:
                    outs.append(None)
                    continue
                if final == "merge":
                    part = ops.merge_runs(wt[b][:fill], tt[b][:fill], offs)
                else:
                    if merge_done is not None:
                        streams[b].wait_event(merge_done)
                    ops.sort_u64(wt[b][:fill], tt[b][:fill])
                    merge_done = torch.cuda.Event()
                    merge_done.record(streams[b])
                    part = wt[b][:fill]
                if check:
                    sp = _u(part)
                    ok_dev &= (sp[1:] >= sp[:-1]).all()
                    edges[q, 0], edges[q, 1] = sp[0], sp[-1]
                    add_checksum(out_acc, part)
                dst, _tier = timed("place", out_spill.place, fill * 8)
                src_alloc = work[b] if part.data_ptr() == wt[b].data_ptr() else tmp[b]
                timed("copy_calls", api.copy, dst, src_alloc, fill * 8, stream=streams[b])
                outs.append(dst)
        barrier()
        t_end = time.time()

        # ---- check: partitions ascending, ordered across all ranks, nothing lost or invented ------------------------
        ok = True
        if check:
            mine = {"ok": bool(ok_dev.item()), "edges": edges.cpu().tolist(), "sizes": sizes,
                    "in": int(in_acc.item()) & (2**64 - 1), "out": int(out_acc.item()) & (2**64 - 1)}
            every = all_gather_obj(mine)
            ok = all(e["ok"] for e in every)
            ok = ok and sum(e["in"] for e in every) % 2**64 == sum(e["out"] for e in every) % 2**64
            ok = ok and sum(sum(e["sizes"]) for e in every) == world * keys_per_rank
            last = None
            for e in every:                                        # rank order = partition order
                for (first, lastk), n in zip(e["edges"], e["sizes"]):
                    if n == 0:
                        continue
                    ok = ok and (last is None or first >= last)
                    last = lastk
        seconds = t_end - t_start
        return {"keys_per_rank": keys_per_rank, "world": world, "n_runs": n_runs, "parts_per_rank": S, "final": final,
                "seconds": seconds, "setup_s": setup_s, "phase1_s": t_runs - t_start, "phase2_s": t_end - t_runs,
                "mkeys_per_s": world * keys_per_rank / seconds / 1e6,
                "bytes_in_hbm": runs_spill.bytes["hbm"] + out_spill.bytes["hbm"],
                "bytes_in_host": runs_spill.bytes["host"] + out_spill.bytes["host"],
                "slices_per_partition": len(sources), "host_blocked_s": {k: round(v, 4) for k, v in host_s.items()},
                "ok": bool(ok)}
    finally:
        torch.cuda.synchronize()
        if dist:
            dist.barrier()                                         # nobody frees a run another rank may still read
        for h in attached:
            h.free()
        if dist:
            dist.barrier()
        for a in work + tmp:
            a.free()
        runs_spill.free()
        out_spill.free()
