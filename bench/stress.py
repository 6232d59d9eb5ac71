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


# One per rank while the allocation clients run (config 5, "strided fp8-cast copies"): a strided fp32 -> block-scaled
# fp8 put into the neighbour's HBM, in a loop, device-timed per launch.  Prints the median microseconds per put.
PUT_LOOP = r"""
import json, os, sys, time, torch, oncilla_b200 as ocm
ocm.init()
torch.cuda.set_device(ocm.device())      # events and synchronize() must look at the GPU the library launches on
rows, elems, pitch = 8192, 4096, 8192
src = ocm.alloc(ocm.OCM_LOCAL_GPU, rows * pitch * 4)
src.local_tensor(torch.float32).normal_()
payload = rows * elems
dst = ocm.alloc(ocm.OCM_REMOTE_GPU, 0, payload + rows * (elems // 32) * 4, flags=ocm.ALLOC_NO_WINDOW)
st = torch.cuda.current_stream()
def step():
    ocm.copy_strided(dst, src, row_bytes=elems * 4, counts=(rows, 1, 1), src_strides=(pitch * 4, 0, 0),
                     dst_strides=(elems, 0, 0), src_dtype=ocm.DT_F32, dst_dtype=ocm.DT_F8E4M3, scale_block=32,
                     scale_offset=payload, stream=st)
for _ in range(5): step()
torch.cuda.synchronize()
stop, us = sys.argv[1], []
t_end = time.time() + float(sys.argv[2])
while not os.path.exists(stop) and time.time() < t_end:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); [step() for _ in range(10)]; e1.record(); torch.cuda.synchronize()
    us.append(e0.elapsed_time(e1) * 100.0)
us.sort()
print("PUT " + json.dumps({"rank": ocm.rank(), "to": dst.remote_rank, "puts": 10 * len(us), "us_p50": us[len(us) // 2],
                           "us_p90": us[len(us) * 9 // 10], "wire_GBps_p50": (payload + rows * (elems // 32) * 4) / us[len(us) // 2] / 1e3}))
dst.free(); src.free(); ocm.tini()
"""


def run_put_loops(cl, ranks, stop_file, seconds):
    import os

    procs = []
    for r in range(ranks):
        env = cl.env(r)
        env["PYTHONPATH"] = str(REPO) + os.pathsep + env.get("PYTHONPATH", "")
        procs.append(subprocess.Popen([sys.executable, "-c", PUT_LOOP, stop_file, str(seconds)], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    return procs


def collect_puts(procs):
    rows = []
    for p in procs:
        out, err = p.communicate(timeout=600)
        for line in out.splitlines():
            if line.startswith("PUT "):
                rows.append(json.loads(line[4:]))
        if p.returncode:
            print(err[-800:], file=sys.stderr)
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--apps", type=int, default=64)
    ap.add_argument("--iters", type=int, default=10000)
    ap.add_argument("--ranks", type=int, default=2)
    ap.add_argument("--kind", type=int, default=8)
    ap.add_argument("--max-bytes", type=int, default=65536)
    ap.add_argument("--gpu", action="store_true", help="daemons own HBM pools (needed for --kind 7)")
    ap.add_argument("--puts", action="store_true", help="one strided fp32->fp8 put loop per rank runs next to the clients")
    ap.add_argument("--gpu-bytes", type=int, default=4 << 30)
    args = ap.parse_args()
    extra = {} if args.gpu else {"OCM_NO_GPU": "1"}
    kw = dict(gpus=list(range(args.ranks)), gpu_bytes=args.gpu_bytes) if args.gpu else {}
    with ocm.LocalCluster(args.ranks, host_bytes=512 << 20, extra_env=extra, **kw) as cl:
        import tempfile

        stop_file = tempfile.mktemp(prefix="ocm_stress_stop_")
        puts_alone = puts_loaded = None
        if args.puts:      # the put loops alone first: what the clients' allocations have to be measured against
            puts_alone = collect_puts(run_put_loops(cl, args.ranks, stop_file, 3.0))
        put_procs = run_put_loops(cl, args.ranks, stop_file, 600.0) if args.puts else []
        if put_procs:
            time.sleep(4.0)                        # imports + warm-up of the put loops
        # how much of the pools the small blocks really cost, sampled while the clients run
        cl.apply_env(0)
        ocm.init()
        peak = {"requested_max": 0, "payload": 0, "granules": 0}
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
        tier = "gpu" if args.kind in (6, 7) else "host"
        while any(p.poll() is None for _, p in procs):
            pay = gran = 0
            for r in range(args.ranks):
                st = ocm.stats(rank=r)
                pay += st[f"pool_{tier}_used"]
                gran += st[f"pool_{tier}_granule_bytes"]
            if gran > peak["granules"]:
                peak["payload"], peak["granules"] = pay, gran
            time.sleep(0.05)
        for path, p in procs:
            out = p.communicate(timeout=1800)[0]
            m = re.search(r"alloc_us p50=([\d.]+) p99=([\d.]+) max=([\d.]+)\s+free_us p50=([\d.]+)", out)
            if p.returncode != 0 or not m:
                failed += 1
                print(out[-500:], file=sys.stderr)
                continue
            res[path].append(tuple(float(x) for x in m.groups()))
        wall = time.time() - t0
        if put_procs:
            open(stop_file, "w").close()
            puts_loaded = collect_puts(put_procs)
        time.sleep(0.5)
        live = [ocm.stats(rank=r)["allocs_live"] for r in range(args.ranks)]
        faults = [ocm.stats(rank=r)["table_faults"] for r in range(args.ranks)]
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
        "live_after": live, "table_faults": faults,
        # every client holds up to 8 blocks of at most max_bytes: requested bytes are bounded by apps * 8 * max_bytes
        "pool_peak": {"tier": tier, "granule_bytes": peak["granules"], "size_class_bytes": peak["payload"],
                      "requested_upper_bound": args.apps * 8 * args.max_bytes,
                      "granules_over_size_classes": round(peak["granules"] / peak["payload"], 2) if peak["payload"] else None},
        "puts_alone": puts_alone, "puts_with_clients": puts_loaded,
        "put_slowdown_p50": (round(sorted(r["us_p50"] for r in puts_loaded)[len(puts_loaded) // 2] /
                                   sorted(r["us_p50"] for r in puts_alone)[len(puts_alone) // 2], 3)
                             if puts_alone and puts_loaded else None)}), flush=True)
    return 1 if failed or any(live) or any(faults) else 0


if __name__ == "__main__":
    sys.exit(main())
