#!/usr/bin/env python
"""C-level latency numbers through the public API (no Python in the timed loops):
the native `ocm_test 4` sweep (blocking ocm_copy_onesided per message, host clock - the reference's own BW
test shape, reference test/ocm_test.c:323-425) and `ocm_stress` allocation latency for remote-GPU pairs.

    python bench/latency_probe.py [--ranks 2]
"""
import argparse
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

import oncilla_b200 as ocm  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--ranks", type=int, default=2)
args = ap.parse_args()
BIN = REPO / "build" / "bin"
with ocm.LocalCluster(args.ranks, gpus=list(range(args.ranks)), gpu_bytes=6 << 30, host_bytes=256 << 20) as cl:
    for argv in (["ocm_test", "4", "1", "1", "7", "300"],            # remote-GPU pair, 64 B .. 1 MiB, 300 iterations
                 ["ocm_stress", "3000", "1048576", "7", "4"],        # alloc/free of <= 1 MiB remote-GPU pairs
                 ["ocm_stress", "2000", "268435456", "7", "2"]):     # ... of <= 256 MiB
        # the blocking-copy sweep runs with and without the resident copy server (H2); the allocation loops on both paths
        for path in (("fast", "server=1"), ("fast", "server=0")) if argv[0] == "ocm_test" else (("fast", ""), ("daemon", "")):
            env = cl.env(0)
            env["OCM_ALLOC_PATH"] = path[0]
            if path[1]:
                env["OCM_SERVER"] = path[1][-1]
                env["OCM_SERVER_TRACE"] = "1"
            r = subprocess.run([str(BIN / argv[0]), *argv[1:]], env=env, capture_output=True, text=True, timeout=600)
            print(f"### {' '.join(argv)} [{' '.join(path).strip()}] rc={r.returncode}")
            print(r.stdout[-3000:])
            print("\n".join(l for l in r.stderr.splitlines() if "copy_server_trace" in l))
            if r.returncode:
                print(r.stderr[-1000:])
