#!/usr/bin/env python
"""Many-client allocation stress (BASELINE.json config 5): APPS processes x ITERS ocm_alloc/ocm_free each,
spread over the ranks of a local cluster, half of them on the fast path and half through the daemon protocol.

    python bench/stress.py [--apps 64] [--iters 10000] [--ranks 2] [--kind 8|7] [--max-bytes 65536]

Prints one JSON line with the latency distribution per path and the final pool state (must be empty).
"""
from __future__ import annotations

import argparse
import json
import re
import subprocess
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

import oncilla_b200 as ocm  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--apps", type=int, default=64)
    ap.add_argument("--iters", type=int, default=10000)
    ap.add_argument("--ranks", type=int, default=2)
    ap.add_argument("--kind", type=int, default=8)
    ap.add_argument("--max-bytes", type=int, default=65536)
    ap.add_argument("--gpu", action="store_true", help="daemons own HBM pools (needed for --kind 7)")
    args = ap.parse_args()
    extra = {} if args.gpu else {"OCM_NO_GPU": "1"}
    kw = dict(gpus=list(range(args.ranks)), gpu_bytes=2 << 30) if args.gpu else {}
    with ocm.LocalCluster(args.ranks, host_bytes=512 << 20, extra_env=extra, **kw) as cl:
        t0 = time.time()
        procs = []
        for i in range(args.apps):
            env = cl.env(i % args.ranks)
            env["OCM_ALLOC_PATH"] = "fast" if i % 2 == 0 else "daemon"
            procs.append((env["OCM_ALLOC_PATH"], subprocess.Popen(
                [str(REPO / "build" / "bin" / "ocm_stress"), str(args.iters), str(args.max_bytes), str(args.kind)],
                env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        res = {"fast": [], "daemon": []}
        failed = 0
        for path, p in procs:
            out = p.communicate(timeout=1800)[0]
            m = re.search(r"alloc_us p50=([\d.]+) p99=([\d.]+) max=([\d.]+)\s+free_us p50=([\d.]+)", out)
            if p.returncode != 0 or not m:
                failed += 1
                print(out[-500:], file=sys.stderr)
                continue
            res[path].append(tuple(float(x) for x in m.groups()))
        wall = time.time() - t0
        cl.apply_env(0)
        ocm.init()
        time.sleep(0.5)
        live = [ocm.stats(rank=r)["allocs_live"] for r in range(args.ranks)]
        ocm.tini()

    def agg(rows, i):
        v = sorted(r[i] for r in rows)
        return round(v[len(v) // 2], 2) if v else None

    print(json.dumps({
        "kind": "stress", "apps": args.apps, "iters": args.iters, "ranks": args.ranks, "alloc_kind": args.kind,
        "failed_apps": failed, "wall_s": round(wall, 2),
        "allocs_per_s": round(args.apps * args.iters / wall),
        "fast": {"apps": len(res["fast"]), "alloc_p50_us": agg(res["fast"], 0), "alloc_p99_us": agg(res["fast"], 1),
                 "free_p50_us": agg(res["fast"], 3)},
        "daemon": {"apps": len(res["daemon"]), "alloc_p50_us": agg(res["daemon"], 0),
                   "alloc_p99_us": agg(res["daemon"], 1), "free_p50_us": agg(res["daemon"], 3)},
        "live_after": live}), flush=True)
    return 1 if failed or any(live) else 0


if __name__ == "__main__":
    sys.exit(main())
