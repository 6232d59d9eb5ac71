"""``python -m oncilla_b200.stat`` - one line per rank of a running cluster: pool sizes, usage, live blocks.

Attaches like any application (``OCM_RANK`` / ``OCM_NS`` select the cluster), asks every
daemon for its counters (``ocm_get_rank_stats`` -> ``MSG_STATS``, forwarded by the local daemon) and leaves.
The reference has no observability beyond ``OCM_VERBOSE`` printf lines.
"""
from __future__ import annotations

import json
import sys

from . import api


def _fmt(n: int) -> str:
    for unit, div in (("GiB", 1 << 30), ("MiB", 1 << 20), ("KiB", 1 << 10)):
        if n >= div:
            return f"{n / div:.1f} {unit}"
    return f"{n} B"


def collect() -> list[dict]:
    rows = []
    for r in range(api.num_ranks()):
        try:
            s = api.stats(r)
            s["rank"] = r
            s["reachable"] = True
        except Exception as e:  # a dead daemon must not hide the others
            s = {"rank": r, "reachable": False, "error": str(e)}
        rows.append(s)
    return rows


def main(argv=None) -> int:
    argv = sys.argv[1:] if argv is None else argv
    api.init()
    try:
        rows = collect()
    finally:
        api.tini()
    if "--json" in argv:
        print(json.dumps(rows))
        return 0
    if "--prometheus" in argv:
        # text exposition format: scrape it from a cron job / node_exporter textfile collector
        print("# TYPE ocm_rank_up gauge")
        for s in rows:
            print(f'ocm_rank_up{{rank="{s["rank"]}"}} {1 if s["reachable"] else 0}')
        for key, help_ in (("pool_gpu_bytes", "HBM pool capacity"), ("pool_gpu_used", "HBM pool bytes in use"),
                           ("pool_host_bytes", "pinned host pool capacity"), ("pool_host_used", "pinned host pool bytes in use"),
                           ("pool_gpu_granule_bytes", "HBM pool bytes taken in granules (blocks + slab remainders)"),
                           ("pool_host_granule_bytes", "host pool bytes taken in granules"),
                           ("table_faults", "table self-checks that found an inconsistency"),
                           ("allocs_live", "live blocks"), ("allocs_total", "allocations since start"),
                           ("frees_total", "frees since start")):
            print(f"# HELP ocm_{key} {help_}\n# TYPE ocm_{key} {'counter' if key.endswith('_total') else 'gauge'}")
            for s in rows:
                if s["reachable"] and key in s:
                    print(f'ocm_{key}{{rank="{s["rank"]}"}} {s[key]}')
        return 0
    print(f"{'rank':>4}  {'HBM pool':>10} {'used':>10}  {'host pool':>10} {'used':>10}  {'live':>6} {'allocs':>8} {'frees':>8} {'faults':>6}")
    for s in rows:
        if not s["reachable"]:
            print(f"{s['rank']:>4}  unreachable: {s['error']}")
            continue
        print(f"{s['rank']:>4}  {_fmt(s['pool_gpu_bytes']):>10} {_fmt(s['pool_gpu_used']):>10}  "
              f"{_fmt(s['pool_host_bytes']):>10} {_fmt(s['pool_host_used']):>10}  "
              f"{s['allocs_live']:>6} {s['allocs_total']:>8} {s['frees_total']:>8} {s.get('table_faults', 0):>6}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
