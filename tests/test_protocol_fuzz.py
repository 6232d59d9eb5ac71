"""Hostile input on every door of a daemon: the peer TCP port, the unix fd server and the POSIX message queue.

The reference trusts its wire completely (fixed 160-byte struct read with no magic, no version, no length check:
reference src/mem.c:62-111, src/pmsg.c:452-475).  Here every message carries magic + version + size; garbage must
be dropped without taking the daemon down or wedging a later, well-formed client.
"""
import ctypes
import json
import os
import random
import socket
import struct

from conftest import BIN, run_app, run_py

import oncilla_b200 as ocm

ALLOC_SNIPPET = r"""
import json, oncilla_b200 as ocm
ocm.init()
res = []
for flags in (0, ocm.ALLOC_SLOW_PATH):
    a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20, flags=flags)
    a.local_tensor()[:] = 3; a.put(); a.local_tensor()[:] = 0; a.get()
    res.append([a.remote_rank, int(a.local_tensor().sum()) == 3 * 4096])
    a.free()
live = [ocm.stats(rank=r)["allocs_live"] for r in range(ocm.num_ranks())]
ocm.tini()
print("RESULT " + json.dumps({"res": res, "live": live}))
"""


def _ports(cl):
    ports = []
    for line in open(cl.nodefile):
        f = line.split()
        if f and not f[0].startswith("#"):
            ports.append(int(f[3]))
    return ports


def _healthy(cl):
    out = run_py(cl, 0, ALLOC_SNIPPET)
    assert out.returncode == 0, out.stdout + out.stderr
    res = json.loads(out.stdout.split("RESULT ", 1)[1])
    assert res["res"] == [[1, True], [1, True]] and res["live"] == [0, 0], res


def test_garbage_on_tcp_uds_and_mq_does_not_hurt_the_daemons():
    rng = random.Random(7)
    with ocm.LocalCluster(2, host_bytes=16 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:
        _healthy(cl)
        # --- peer TCP ports: random bytes, short writes, a huge blob, half-open connections -------------------------
        for port in _ports(cl):
            for n in (1, 7, 191, 192, 193, 4096, 1 << 20):
                s = socket.create_connection(("127.0.0.1", port), timeout=5)
                try:
                    s.sendall(bytes(rng.randrange(256) for _ in range(min(n, 65536))) * max(1, n // 65536))
                except OSError:
                    pass                                  # the daemon may hang up on us: that is fine
                s.close()
            idle = [socket.create_connection(("127.0.0.1", port), timeout=5) for _ in range(8)]   # connect and say nothing
            # a frame with the right size but a wrong magic / version
            s = socket.create_connection(("127.0.0.1", port), timeout=5)
            s.sendall(struct.pack("<IHH", 0xDEADBEEF, 999, 5) + bytes(184))
            s.close()
            for c in idle:
                c.close()
        _healthy(cl)
        # --- unix fd server (abstract socket): garbage and an unknown request ----------------------------------------
        for rank in range(2):
            name = "\0ocm_" + cl.ns + "_uds_r" + str(rank)
            for payload in (b"x", bytes(rng.randrange(256) for _ in range(64)), struct.pack("<IIii", 0x4F434D55, 77, os.getpid(), 0)):
                s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                s.settimeout(5)
                try:
                    s.connect(name)
                    s.sendall(payload)
                    try:
                        s.recv(64)
                    except OSError:
                        pass
                except OSError:
                    pass
                s.close()
        _healthy(cl)
        # --- the daemons' message queues: wrong sizes and wrong magic ------------------------------------------------------
        librt = ctypes.CDLL("librt.so.1", use_errno=True)
        librt.mq_open.restype = ctypes.c_int
        for rank in range(2):
            qname = ("/ocm_mq_" + cl.ns + "_daemon" + (f"_r{rank}" if rank else "")).encode()
            q = librt.mq_open(qname, os.O_WRONLY | os.O_NONBLOCK)
            assert q >= 0, f"mq_open {qname}: errno {ctypes.get_errno()}"
            for n in (0, 1, 100, 192):
                buf = bytes(rng.randrange(256) for _ in range(n))
                librt.mq_send(q, buf, ctypes.c_size_t(n), 0)          # may fail with EAGAIN when full: fine
            librt.mq_close(q)
        _healthy(cl)
        # native client still fine, and the daemons are the same processes as at the start
        r = run_app(cl, 1, [BIN / "ocm_test", "3", "1", "1", "1"])
        assert r.returncode == 0 and "pass: OCM test completed successfully" in r.stdout, r.stdout + r.stderr
        assert all(p.poll() is None for p in cl.procs.values())


def _msg_version():
    import re
    from conftest import REPO

    return int(re.search(r"kMsgVersion\s*=\s*(\d+)", (REPO / "src" / "common" / "msg.hpp").read_text()).group(1))


def _frame(mtype, payload, status=1, rank=1, seq=77):
    """A well-formed 192-byte Message (src/common/msg.hpp): header (32 B) + 160-byte union."""
    hdr = struct.pack("<IHHHHiiIi", 0x4F434D32, _msg_version(), mtype, status, 0, os.getpid(), rank, seq, 0)
    hdr += bytes(32 - len(hdr))
    return hdr + payload + bytes(160 - len(payload))


def _exchange(port, frame, expect_reply=True):
    s = socket.create_connection(("127.0.0.1", port), timeout=5)
    s.settimeout(3)
    s.sendall(frame)
    data = b""
    try:
        while expect_reply and len(data) < 192:
            chunk = s.recv(192 - len(data))
            if not chunk:
                break
            data += chunk
    except OSError:
        pass
    s.close()
    return data


def test_huge_sizes_and_bad_tiers_on_the_wire_are_refused():
    """Well-formed frames with sizes near 2^64 and out-of-range tiers: `used + bytes` and
    `(bytes + granule - 1) / granule` must not wrap, and a wire tier must never index past the cluster table."""
    MSG_REQ_ALLOC, MSG_DO_ALLOC, MSG_DO_FREE, MSG_NOTE_FREE = 5, 6, 8, 16
    ALLOC_MEM_REMOTE_HOST, TIER_HOST, OCM_POLICY_DEFAULT = 6, 1, 0
    with ocm.LocalCluster(2, host_bytes=16 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:
        _healthy(cl)
        p0, p1 = _ports(cl)
        # something must be reserved first: the wrapping sum is only accepted when used > 0
        out = run_py(cl, 0, r"""
import oncilla_b200 as ocm
ocm.init()
keep = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20)
for huge in (2**64 - 1, 2**64 - 4096, 2**63, (16 << 20) + 1):
    for flags in (0, ocm.ALLOC_SLOW_PATH):
        try:
            ocm.alloc(ocm.OCM_REMOTE_HOST, 0, huge, flags=flags | ocm.ALLOC_NO_WINDOW)
            print("ACCEPTED", huge, flags)
        except ocm.OcmError:
            pass
keep.free()
ocm.tini()
print("DONE")
""")
        assert out.returncode == 0 and "DONE" in out.stdout and "ACCEPTED" not in out.stdout, out.stdout + out.stderr
        for huge in (2**64 - 1, 2**64 - 4096, 2**63):
            # placement request straight to rank 0
            req = struct.pack("<iiQIIII", 1, -1, huge, ALLOC_MEM_REMOTE_HOST, OCM_POLICY_DEFAULT, 0, 0)
            rep = _exchange(p0, _frame(MSG_REQ_ALLOC, req))
            assert len(rep) == 192 and struct.unpack_from("<i", rep, 24)[0] != 0, "rank 0 placed a 2^64-byte block"
            # fulfilment request straight to the owner rank
            alloc = struct.pack("<iiQIIQQQIIII", 0, 1, 0, ALLOC_MEM_REMOTE_HOST, TIER_HOST, huge, 0, 0, 0, 0, 0, 0)
            rep = _exchange(p1, _frame(MSG_DO_ALLOC, alloc, rank=0))
            assert len(rep) == 192 and struct.unpack_from("<i", rep, 24)[0] != 0, "owner carved a 2^64-byte block"
        for tier in (2, 7, 99, 0xFFFFFFFF):
            alloc = struct.pack("<iiQIIQQQIIII", 0, 1, 1, ALLOC_MEM_REMOTE_HOST, tier, 4096, 0, 0, 0, 0, 0, 0)
            _exchange(p0, _frame(MSG_NOTE_FREE, alloc, status=0), expect_reply=False)
            rep = _exchange(p1, _frame(MSG_DO_FREE, alloc))
            assert len(rep) == 192 and struct.unpack_from("<i", rep, 24)[0] != 0
            rep = _exchange(p1, _frame(MSG_DO_ALLOC, alloc))
            assert len(rep) == 192 and struct.unpack_from("<i", rep, 24)[0] != 0
        _healthy(cl)
        assert all(p.poll() is None for p in cl.procs.values())
