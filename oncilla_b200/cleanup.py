"""``python -m oncilla_b200.cleanup [--ns TAG] [--dry-run]`` - remove IPC objects that dead daemons left behind.

A daemon that exits normally unlinks its shared-memory segments and message queue; one that is killed with SIGKILL
(or the box's OOM killer) cannot.  A restart in the same namespace replaces them, but abandoned namespaces stay in
``/dev/shm`` (the pinned host pool can be gigabytes).  This sweeps ``/dev/shm/ocm_*tbl_r<rank>``: when the daemon pid
recorded in the table header no longer exists, the rank's table, its host pool, rank 0's cluster table and the
daemon's message queue are removed.  Counterpart of the reference's ``pmsg_cleanup`` (reference src/pmsg.c:373-412,
message queues only, swept by pid_max).  Live clusters are never touched.
"""
from __future__ import annotations

import argparse
import ctypes
import os
import re
import struct
import sys
from pathlib import Path

SHM = Path("/dev/shm")
TABLE_MAGIC = 0x4F434D5441424C45          # "OCMTABLE", src/common/tables.hpp
_TBL = re.compile(r"^ocm_(?:(?P<ns>.+)_)?tbl_r(?P<rank>\d+)$")


def _daemon_pid(path: Path):
    """pid stored in a table header: magic u64, version u32, rank i32, n_ranks i32, daemon_pid i32."""
    try:
        with open(path, "rb") as f:
            head = f.read(24)
        magic, _ver, _rank, _n, pid = struct.unpack("<QIiii", head)
        return pid if magic == TABLE_MAGIC else None
    except (OSError, struct.error):
        return None


def _alive(pid: int) -> bool:
    if pid <= 0:
        return False
    try:
        os.kill(pid, 0)
        return True
    except ProcessLookupError:
        return False
    except PermissionError:
        return True


def sweep(ns: str | None = None, dry_run: bool = False) -> list[str]:
    """Remove what dead daemons left (all namespaces, or only ``ns``).  Returns the names removed."""
    removed: list[str] = []
    if not SHM.is_dir():
        return removed
    try:
        librt = ctypes.CDLL("librt.so.1", use_errno=True)
    except OSError:
        librt = None
    for p in sorted(SHM.iterdir()):
        m = _TBL.match(p.name)
        if not m or (ns is not None and (m.group("ns") or "") != ns):
            continue
        pid = _daemon_pid(p)
        if pid is None or _alive(pid):
            continue
        tag, rank = m.group("ns"), int(m.group("rank"))
        prefix = f"ocm_{tag}_" if tag else "ocm_"
        victims = [p, SHM / f"{prefix}host_r{rank}"]
        if rank == 0:
            victims.append(SHM / f"{prefix}cluster")
        for v in victims:
            if v.exists():
                removed.append(v.name)
                if not dry_run:
                    try:
                        v.unlink()
                    except OSError:
                        pass
        mq = "/ocm_mq_" + (f"{tag}_" if tag else "") + "daemon" + (f"_r{rank}" if rank else "")
        if librt is not None and not dry_run and librt.mq_unlink(mq.encode()) == 0:
            removed.append(mq)
    return removed


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("--ns", default=None, help="only this OCM_NS namespace (default: every namespace)")
    ap.add_argument("--dry-run", action="store_true", help="list what would be removed")
    args = ap.parse_args(argv)
    names = sweep(args.ns, args.dry_run)
    for n in names:
        print(("would remove " if args.dry_run else "removed ") + n)
    print(f"{len(names)} object(s)")
    return 0


if __name__ == "__main__":
    sys.exit(main())
