"""Checkpoint / restore of allocation contents.

The reference has no persistence of any kind: remote memory lives exactly as long as the daemons.
This module writes the payload of a set of named allocations to one file and puts it back later - into new
allocations, possibly on another placement, after a daemon or cluster restart.

    from oncilla_b200 import checkpoint as ckpt
    ckpt.save("state.ocm", {"table": a, "keys": b})
    ...restart...
    a2, b2 = ocm.alloc(...), ocm.alloc(...)
    ckpt.load("state.ocm", {"table": a2, "keys": b2})

Payload of an allocation = the far side of a pair (peer HBM / peer pinned host), else the local memory.  Data moves
in 64 MiB pieces through the tensor views of the binding (device <-> host copies for HBM); every entry carries a
CRC-32 that ``load`` verifies.

File layout: 8-byte magic ``OCMCKPT1``, u64 header length, JSON header, payloads.  The first MiB is reserved for the
header (``data_start`` = 1 MiB); entry offsets are relative to it and 4 KiB-aligned.
"""
from __future__ import annotations

import json
import struct
import zlib

from . import _native as N
from .api import Allocation

MAGIC = b"OCMCKPT1"
PIECE = 64 << 20
DATA_START = 1 << 20


class CheckpointError(RuntimeError):
    pass


def _payload(a: Allocation):
    """The uint8 tensor that holds what the allocation stores."""
    if a.kind in (N.OCM_STRIPED_GPU, N.OCM_REPLICATED_GPU):
        raise CheckpointError("multi-slab kinds are not supported: checkpoint the window or the slabs' owners")
    if a.is_remote and a.remote_ptr:
        return a.remote_tensor()
    return a.local_tensor()


def save(path: str, allocs: dict[str, Allocation]) -> dict:
    """Write the payload of every allocation to ``path``.  Returns the header that was written."""
    entries, offset = [], 0
    for name, a in allocs.items():
        n = int(_payload(a).numel())
        entries.append({"name": str(name), "kind": int(a.kind), "bytes": n, "offset": offset, "crc32": 0})
        offset += (n + 4095) & ~4095
    with open(path, "wb") as f:
        for e, a in zip(entries, allocs.values()):
            t = _payload(a)
            f.seek(DATA_START + e["offset"])
            crc = 0
            for o in range(0, e["bytes"], PIECE):
                buf = t[o : o + PIECE].cpu().contiguous().numpy().tobytes()    # .cpu(): D2H for HBM, no-op for host
                crc = zlib.crc32(buf, crc)
                f.write(buf)
            e["crc32"] = crc
        header = {"version": 1, "data_start": DATA_START, "entries": entries}
        blob = json.dumps(header).encode()
        if 16 + len(blob) > DATA_START:
            raise CheckpointError("too many entries for the reserved header area")
        f.seek(0)
        f.write(MAGIC + struct.pack("<Q", len(blob)) + blob)
    return header


def read_header(path: str) -> dict:
    """The JSON header of a checkpoint (names, sizes, CRCs) without touching the payloads."""
    with open(path, "rb") as f:
        return _read_header(f, path)


def _read_header(f, path: str) -> dict:
    head = f.read(16)
    if len(head) < 16 or head[:8] != MAGIC:
        raise CheckpointError(f"{path}: not an OCM checkpoint")
    (hlen,) = struct.unpack("<Q", head[8:])
    if hlen == 0 or 16 + hlen > DATA_START:
        raise CheckpointError(f"{path}: implausible header length {hlen}")
    try:
        header = json.loads(f.read(hlen))
        assert header["version"] == 1 and isinstance(header["entries"], list) and header["data_start"] >= 16 + hlen
    except (ValueError, KeyError, AssertionError) as e:
        raise CheckpointError(f"{path}: corrupt header ({e})") from None
    return header


def load(path: str, allocs: dict[str, Allocation], *, strict: bool = True) -> None:
    """Restore the payloads named in ``allocs`` from ``path``.  Sizes must match; CRCs are verified.

    With ``strict=False`` names missing from the file are skipped instead of raising."""
    import numpy as np
    import torch

    with open(path, "rb") as f:
        header = _read_header(f, path)
        by_name = {e["name"]: e for e in header["entries"]}
        for name, a in allocs.items():
            e = by_name.get(str(name))
            if e is None:
                if strict:
                    raise CheckpointError(f"{path}: no entry named {name!r}")
                continue
            t = _payload(a)
            if int(t.numel()) != e["bytes"]:
                raise CheckpointError(f"{name}: checkpoint holds {e['bytes']} bytes, the allocation has {int(t.numel())}")
            f.seek(header["data_start"] + e["offset"])
            crc = 0
            for o in range(0, e["bytes"], PIECE):
                want = min(PIECE, e["bytes"] - o)
                buf = f.read(want)
                if len(buf) != want:
                    raise CheckpointError(f"{name}: file truncated")
                crc = zlib.crc32(buf, crc)
                t[o : o + want].copy_(torch.from_numpy(np.frombuffer(buf, dtype=np.uint8).copy()))
            if crc != e["crc32"]:
                raise CheckpointError(f"{name}: CRC mismatch (header {e['crc32']:#x}, data {crc:#x})")
    if torch.cuda.is_available():
        torch.cuda.synchronize()
