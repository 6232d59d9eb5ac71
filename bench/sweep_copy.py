#!/usr/bin/env python
"""Bandwidth sweeps (device-timed): message size x kernel tuning, for put / get / local copy, against the
library baselines cudaMemcpyAsync (peer or local) - BASELINE.md section 4.

    python bench/sweep_copy.py [--ranks 1|2] [--quick] [--out gpurun_out/sweep.jsonl]

With --ranks 2 the far side lives in GPU 1's pool: puts/gets cross NVLink.  One process (rank 0) drives.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

import torch  # noqa: E402

import oncilla_b200 as ocm  # noqa: E402
from oncilla_b200 import ops  # noqa: E402


def timed(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 1e3 / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ranks", type=int, default=min(torch.cuda.device_count(), 2))
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--out", default=str(REPO / "gpurun_out" / "sweep.jsonl"))
    args = ap.parse_args()
    Path(args.out).parent.mkdir(parents=True, exist_ok=True)
    out = open(args.out, "a")

    def emit(**kw):
        line = json.dumps(kw)
        print(line, flush=True)
        out.write(line + "\n")
        out.flush()

    torch.cuda.set_device(0)
    big = 1 << 30
    with ocm.LocalCluster(args.ranks, gpus=list(range(args.ranks)), gpu_bytes=3 * big, host_bytes=big + (64 << 20)) as cl:
        cl.apply_env(0)
        ocm.init()
        pair = ocm.alloc(ocm.OCM_REMOTE_GPU, big, big)
        loc = ocm.alloc(ocm.OCM_LOCAL_GPU, big)
        hst = ocm.alloc(ocm.OCM_LOCAL_HOST, big)
        win, far, src, hp = pair.local_tensor(), pair.remote_tensor(), loc.local_tensor(), hst.local_tensor()
        where = "peer" if pair.remote_rank != 0 else "local"
        emit(kind="setup", ranks=args.ranks, far_side=where, remote_rank=pair.remote_rank)

        # --- 1. tuning sweep at one large size --------------------------------------------------------
        n = 256 << 20
        grid = [(c, a, s, t) for c in (8, 16, 32, 64, 148) for a in (1, 2, 4, 8) for s in (2, 3, 4) for t in (8192, 16384, 32768)
                if a * s * t <= 224 << 10]
        if args.quick:
            grid = [(c, 4, 3, 16384) for c in (16, 32, 64, 148)] + [(148, 8, 3, 8192), (148, 2, 3, 32768), (64, 4, 4, 8192)]
        for c, a, s, t in grid:
            for name, d, sp in (("put", far, win), ("get", win, far)):
                sec = timed(lambda: ops.copy(d, sp, n, mode=1, ctas=c, agents=a, stages=s, tile_bytes=t), 5, 2)
                emit(kind="tune", op=name, far_side=where, bytes=n, ctas=c, agents=a, stages=s, tile=t, GBps=round(n / sec / 1e9, 1))
        for c in (8, 16, 32, 64, 148, 296, 592):
            for name, d, sp in (("put", far, win), ("get", win, far)):
                sec = timed(lambda: ops.copy(d, sp, n, mode=2, ctas=c), 5, 2)
                emit(kind="tune_vec", op=name, far_side=where, bytes=n, ctas=c, GBps=round(n / sec / 1e9, 1))

        # --- 2. size sweep 4 KB .. 1 GB with default tuning, vs cudaMemcpyAsync -------------------------
        sizes = [4 << 10 << (2 * i) for i in range(10)]      # 4K, 16K, ... 1G
        for n in sizes:
            it = 50 if n <= (1 << 20) else 10 if n <= (64 << 20) else 4
            row = {"kind": "size", "far_side": where, "bytes": n}
            row["put_GBps"] = round(n / timed(lambda: ops.copy(far, win, n), it) / 1e9, 2)
            row["get_GBps"] = round(n / timed(lambda: ops.copy(win, far, n), it) / 1e9, 2)
            row["put_dual_GBps"] = round(n / timed(lambda: ops.copy(far, src, n, dst2=win), it) / 1e9, 2)
            row["h2far_zero_copy_GBps"] = round(n / timed(lambda: ops.copy(far, hp, n, mode=1), it) / 1e9, 2) if n >= (1 << 16) else None
            row["memcpy_put_GBps"] = round(n / timed(lambda: far[:n].copy_(win[:n], non_blocking=True), it) / 1e9, 2)
            row["memcpy_get_GBps"] = round(n / timed(lambda: win[:n].copy_(far[:n], non_blocking=True), it) / 1e9, 2)
            # the public API path (ocm_copy_onesided_async on the current stream)
            st = torch.cuda.current_stream()
            row["api_put_GBps"] = round(n / timed(lambda: pair.put(n, stream=st), it) / 1e9, 2)
            row["api_put_us"] = round(timed(lambda: pair.put(n, stream=st), it) * 1e6, 2)
            emit(**row)
        for a in (pair, loc, hst):
            a.free()
        ocm.tini()


if __name__ == "__main__":
    main()
