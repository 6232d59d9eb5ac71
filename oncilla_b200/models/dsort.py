"""Distributed sample sort of 64-bit keys: every rank sorts, exchanges and sorts again (K8 pipeline).

One application per GPU (``TorchrunCluster``).  ``torch.distributed`` carries only the control plane (samples,
counts, 32-byte share tokens); every key moves through OCM:

  1. local radix sort of this rank's keys (K8)
  2. regular samples -> splitters (same on every rank)
  3. each rank allocates its receive block in ITS OWN pool, shares it (``ocm_share``) and attaches to the
     blocks of all peers (``ocm_attach``)
  4. all-to-all exchange: one one-sided put per destination straight into the peer's block at the offset the
     count matrix dictates (K1 over NVLink, signal on the destination block)
  5. each rank waits until its block's completion signal has counted every sender (the EXTOLL completer
     notification of the reference, src/extoll.c:145-170, as a counter) and radix-sorts what it received

After step 5 rank r holds the r-th slice of the global order.  Keys are < 2**62 (torch compares signed).
"""
from __future__ import annotations

import time

from .. import _native as N
from .. import api, ops


def _dist():
    import torch.distributed as dist

    return dist if dist.is_available() and dist.is_initialized() else None


def run(keys_per_rank: int = 1 << 22, check: bool = True, seed: int = 0xD5EED, oversample: int = 64) -> dict:
    import torch

    dist = _dist()
    rank = dist.get_rank() if dist else 0
    world = dist.get_world_size() if dist else 1
    dev = torch.device(f"cuda:{api.device()}")
    n = int(keys_per_rank)

    def all_gather(t):
        if not dist:
            return [t]
        out = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(out, t)
        return out

    def barrier():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()

    work = api.alloc(N.OCM_LOCAL_GPU, n * 8, flags=N.ALLOC_NO_ZERO)
    tmp = api.alloc(N.OCM_LOCAL_GPU, n * 8, flags=N.ALLOC_NO_ZERO)
    wt, tt = work.local_tensor(torch.int64), tmp.local_tensor(torch.int64)
    g = torch.Generator(device=dev)
    g.manual_seed(seed + rank)
    wt.copy_(torch.randint(0, 2**62, (n,), dtype=torch.int64, device=dev, generator=g))
    in_sum = ops.checksum(wt) if check else 0
    recv = rtmp = None
    peers = []
    finished = False
    try:
        barrier()
        t0 = time.perf_counter()
        # 1. local sort
        ops.sort_u64(wt, tt)
        # 2. splitters from regular samples of every rank's sorted run
        s = oversample * world
        sample = wt[torch.arange(s, device=dev) * (n // s) + n // (2 * s)].contiguous()
        allsamp = torch.sort(torch.cat(all_gather(sample))).values
        splitters = allsamp[torch.arange(1, world, device=dev) * s] if world > 1 else allsamp[:0]
        bounds = [0] + torch.searchsorted(wt, splitters).tolist() + [n]
        counts = torch.tensor([bounds[p + 1] - bounds[p] for p in range(world)], dtype=torch.int64, device=dev)
        cmat = torch.stack(all_gather(counts)).cpu()          # cmat[src][dst]
        recv_n = int(cmat[:, rank].sum())
        senders = int((cmat[:, rank] > 0).sum())
        t1 = time.perf_counter()
        # 3. receive block in my own pool, opened by everybody
        recv = api.alloc(N.OCM_LOCAL_GPU, max(recv_n, 1) * 8, flags=N.ALLOC_NO_ZERO)
        rtmp = api.alloc(N.OCM_LOCAL_GPU, max(recv_n, 1) * 8, flags=N.ALLOC_NO_ZERO)
        tok = torch.frombuffer(bytearray(recv.share()), dtype=torch.uint8).to(dev)
        tokens = [bytes(t.cpu().numpy().tobytes()) for t in all_gather(tok)]
        peers = [api.attach(t) for t in tokens]
        # 4. one put per destination, farthest neighbour last so that the ranks do not all hit the same peer
        st = torch.cuda.Stream(device=dev)
        st.wait_stream(torch.cuda.current_stream())
        for k in range(world):
            dst = (rank + 1 + k) % world
            cnt = int(cmat[rank, dst])
            if cnt:
                off = int(cmat[:rank, dst].sum())
                api.copy(peers[dst], work, cnt * 8, src_offset=bounds[dst] * 8, dest_offset=off * 8, stream=st)
        # 5. wait for every sender's signal on my block, then sort what arrived
        recv.wait_signal(senders, timeout_ms=20000)
        t2 = time.perf_counter()
        rt = recv.local_tensor(torch.int64)[:recv_n]
        if recv_n:
            ops.sort_u64(rt, rtmp.local_tensor(torch.int64)[:recv_n])
        st.synchronize()
        barrier()
        t3 = time.perf_counter()

        ok = True
        if check:
            out_sum = ops.checksum(rt) if recv_n else 0
            lo = int(rt[0]) if recv_n else -1
            hi = int(rt[-1]) if recv_n else -1
            ok = recv_n < 2 or bool((rt[1:] >= rt[:-1]).all())
            stats = torch.tensor([_wrap(in_sum), _wrap(out_sum), recv_n, lo, hi], dtype=torch.int64, device=dev)
            allst = torch.stack(all_gather(stats)).cpu()
            ok = ok and int(allst[:, 0].sum()) == int(allst[:, 1].sum()) and int(allst[:, 2].sum()) == n * world
            last = -1
            for r in range(world):                            # slices are ordered across ranks
                if int(allst[r, 2]):
                    ok = ok and int(allst[r, 3]) >= last
                    last = int(allst[r, 4])
            ok_t = torch.tensor([1 if ok else 0], device=dev)
            if dist:
                dist.all_reduce(ok_t, op=dist.ReduceOp.MIN)
            ok = bool(int(ok_t))
        el = torch.tensor([t3 - t0, t1 - t0, t2 - t1, t3 - t2], dtype=torch.float64, device=dev)
        if dist:
            dist.all_reduce(el, op=dist.ReduceOp.MAX)
        total, sort1, exch, sort2 = el.tolist()
        finished = True
        return {"world": world, "keys_per_rank": n, "total_keys": n * world, "seconds": total,
                "local_sort_s": sort1, "exchange_s": exch, "final_sort_s": sort2,
                "mkeys_per_s": n * world / total / 1e6, "recv_keys": recv_n,
                "imbalance": recv_n * world / float(n * world), "ok": ok}
    finally:
        for p in peers:
            p.free()
        if finished:                                          # (a failing rank must not wait for the others)
            barrier()                                         # nobody still writes into a block that goes away
        for a in (recv, rtmp, work, tmp):
            if a is not None:
                a.free()


def _wrap(v: int) -> int:
    """A 64-bit checksum as a signed int64 (sums of these wrap like the unsigned ones)."""
    v &= (1 << 64) - 1
    return v - (1 << 64) if v >= 1 << 63 else v
