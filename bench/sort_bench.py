#!/usr/bin/env python
"""K8 radix sort on one GPU: device time per size against torch.sort (CUB pairs sort) and a check against it.

    python bench/sort_bench.py [--log2 20 22 24 26] [--iters 5]
Prints one JSON line per size.  OCM_SORT_IPT=16 selects the wide scatter tile.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402

from oncilla_b200 import ops  # noqa: E402


def timed(fn, iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e30
    for _ in range(iters):
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log2", type=int, nargs="+", default=[20, 22, 24, 26])
    ap.add_argument("--iters", type=int, default=5)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    for lg in args.log2:
        n = 1 << lg
        src = torch.randint(0, 2**62, (n,), dtype=torch.int64, device=dev)
        keys, tmp = src.clone(), torch.empty_like(src)
        ops.sort_u64(keys, tmp)                               # warm-up (scratch allocation) + check
        ref = torch.sort(src).values
        ok = bool(torch.equal(keys, ref))
        odd = src[1 : n // 2 + 1].clone()                     # odd length / 8-byte-aligned base pointer path
        view = src.clone()[1 : n // 2 + 1]
        ops.sort_u64(view, tmp[: view.numel()])
        ok_odd = bool(torch.equal(view, torch.sort(odd).values))

        def ours():
            keys.copy_(src)
            ops.sort_u64(keys, tmp)

        def cub():
            torch.sort(src)

        def copy_only():
            keys.copy_(src)

        t_copy = timed(copy_only, args.iters)
        t_ours = timed(ours, args.iters) - t_copy
        t_cub = timed(cub, args.iters)
        print(json.dumps({"log2_keys": lg, "ours_ms": round(t_ours, 4), "torch_sort_ms": round(t_cub, 4),
                          "ours_Gkeys_s": round(n / t_ours / 1e6, 2), "bytes_moved_GBps": round(24 * 8 * n / t_ours / 1e6, 1),
                          "ok": ok, "ok_unaligned": ok_odd}), flush=True)


if __name__ == "__main__":
    main()
