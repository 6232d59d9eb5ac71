"""Daemon + libocm integration on CPU: several ranks on one host, full protocol, no GPU.

Models the reference's end-to-end program (reference test/ocm_test.c) and adds what it never tested
(SURVEY section 4): data verification, both allocation paths, placement, accounting, double free,
app-death reclamation, daemon-death detection, concurrent clients.
"""
import json
import os
import signal
import subprocess
import sys
import time

import pytest

import oncilla_b200 as ocm
from conftest import BIN, REPO, run_app, run_py


def _ok(r):
    assert r.returncode == 0, f"rc={r.returncode}\n{r.stdout}\n{r.stderr}"
    return r.stdout


@pytest.mark.parametrize("path", ["fast", "daemon"])
@pytest.mark.parametrize("test,args", [(1, ["1", "2", "1"]), (1, ["1", "2", "3"]), (1, ["1", "2", "8"]),
                                       (2, ["1", "1", "3"]), (3, ["1", "1", "1"]), (4, ["0.25", "0.25", "3", "2"]),
                                       (5, ["1", "1", "1"]), (5, ["1", "1", "8"])])
def test_ocm_test_program(cluster2, path, test, args):
    out = _ok(run_app(cluster2, 1, [BIN / "ocm_test", test, *args], env_extra={"OCM_ALLOC_PATH": path}))
    assert "pass: OCM test completed successfully" in out
    assert "FAIL" not in out


def test_example_hello_ocm(cluster2):
    """examples/hello_ocm.c: public header + libocm only, the reference's call sequence."""
    out = _ok(run_app(cluster2, 0, [BIN / "hello_ocm"]))
    assert "pass: hello_ocm" in out and "far side 1048576 bytes on rank 1" in out
    out = _ok(run_app(cluster2, 1, [BIN / "hello_ocm", "3"]))          # the reference's OCM_REMOTE_RDMA kind
    assert "pass: hello_ocm" in out and "on rank 0" in out


@pytest.mark.parametrize("path", ["fast", "daemon"])
def test_one_application_many_threads(cluster2, path):
    """tests/cpp/ocm_mt.c: 8 threads of one process allocate, copy both ways, put into one shared block and free
    concurrently; every byte is verified and the shared block's signal counts every copy exactly once."""
    out = _ok(run_app(cluster2, 0, [BIN / "ocm_mt", "8", "120"], env_extra={"OCM_ALLOC_PATH": path}, timeout=300))
    assert "pass: ocm_mt 8 threads x 120 iterations, 960 shared signals" in out and "FAIL" not in out


def test_example_cxx_raii_layer(cluster2):
    """examples/raii_ocm.cpp: include/oncillamem.hpp (Session / Allocation / exceptions) over the C ABI."""
    out = _ok(run_app(cluster2, 0, [BIN / "raii_ocm"]))
    assert "pass: raii_ocm" in out and "FAIL" not in out
    probe = _ok(run_py(cluster2, 0, "import oncilla_b200 as ocm; ocm.init(); print('LIVE', [ocm.stats(rank=r)['allocs_live'] for r in (0, 1)]); ocm.tini()"))
    assert "LIVE [0, 0]" in probe                                       # the destructors released everything


def test_stat_tool(cluster2):
    """python -m oncilla_b200.stat: per-rank pool counters through MSG_STATS."""
    env = cluster2.env(0)
    env["PYTHONPATH"] = str(REPO) + os.pathsep + env.get("PYTHONPATH", "")
    r = subprocess.run([sys.executable, "-m", "oncilla_b200.stat", "--json"], env=env, capture_output=True, text=True,
                       timeout=60)
    assert r.returncode == 0, r.stderr
    rows = json.loads(r.stdout.strip().splitlines()[-1])
    assert [x["rank"] for x in rows] == [0, 1] and all(x["reachable"] for x in rows)
    assert all(x["pool_host_bytes"] == 32 << 20 for x in rows)
    r = subprocess.run([sys.executable, "-m", "oncilla_b200.stat"], env=env, capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "host pool" in r.stdout and "32.0 MiB" in r.stdout
    r = subprocess.run([sys.executable, "-m", "oncilla_b200.stat", "--prometheus"], env=env, capture_output=True,
                       text=True, timeout=60)
    assert r.returncode == 0 and 'ocm_rank_up{rank="1"} 1' in r.stdout
    assert f'ocm_pool_host_bytes{{rank="0"}} {32 << 20}' in r.stdout and "# TYPE ocm_allocs_total counter" in r.stdout


API_SNIPPET = r"""
import json, torch, ctypes
import oncilla_b200 as ocm
ocm.init()
res = {"rank": ocm.rank(), "n": ocm.num_ranks(), "dev": ocm.device()}
# local host
h = ocm.alloc(ocm.OCM_LOCAL_HOST, 1 << 20)
res["host_kind"] = h.kind; res["host_remote"] = h.is_remote; res["host_rsz"] = h.remote_bytes
t = h.local_tensor()
res["host_zero"] = int(t.sum())
t[:] = torch.arange(1 << 20, dtype=torch.int64).to(torch.uint8)
# remote pair, legacy name -> pinned host on the neighbour
r = ocm.alloc(ocm.OCM_REMOTE_RDMA, 1 << 20, 2 << 20)
res["pair_kind"] = r.kind; res["pair_remote"] = r.is_remote; res["pair_rank"] = r.remote_rank
res["pair_rsz"] = r.remote_bytes; res["far_zero"] = int(r.remote_tensor().sum())
ocm.copy(r, h, 1 << 20)                       # host -> window + far side
res["sig"] = r.signal()
res["far_eq"] = bool(torch.equal(r.remote_tensor()[: 1 << 20], t))
res["win_eq"] = bool(torch.equal(r.local_tensor(), t))
# one-sided at offsets: window[4096:8192] -> far[1 MiB + 100 : ...]
r.put(4096, src_offset=4096, dest_offset=(1 << 20) + 100)
res["off_eq"] = bool(torch.equal(r.remote_tensor()[(1 << 20) + 100:(1 << 20) + 100 + 4096], t[4096:8192]))
res["sig2"] = r.signal()
# read back through a second host allocation (op_flag=0 flips the direction)
h2 = ocm.alloc(ocm.OCM_LOCAL_HOST, 1 << 20)
ocm.copy(r, h2, 1 << 20, op_flag=0)
res["read_eq"] = bool(torch.equal(h2.local_tensor(), t))
# two-hop with distinct offsets: host[0:1024] -> window[2048:3072]; window[512:1536] -> far[0:1024]
before = r.local_tensor()[512:1536].clone()
ocm.copy(r, h, 1024, src_offset=0, dest_offset=2048, src_offset_2=512, dest_offset_2=0)
res["hop1"] = bool(torch.equal(r.local_tensor()[2048:3072], t[:1024]))
res["hop2"] = bool(torch.equal(r.remote_tensor()[:1024], before))
# pair -> pair third-party copy
r2 = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20, policy=ocm.POLICY_SELF)
res["self_rank"] = r2.remote_rank
ocm.copy(r2, r, 1 << 20)
res["p2p_eq"] = bool(torch.equal(r2.remote_tensor(), r.remote_tensor()[: 1 << 20]))
# errors
try:
    r.put(10, dest_offset=(2 << 20))
    res["range_err"] = False
except ocm.OcmError as e:
    res["range_err"] = e.err == 5
st = ocm.stats()
res["live"] = st["allocs_live"]; res["fast"] = st["fast_path_allocs"]; res["slow"] = st["slow_path_allocs"]
rs = ocm.stats(rank=r.remote_rank)
res["remote_pool_used"] = rs["pool_host_used"]
for a in (h, h2, r, r2): a.free()
try:
    h.free(); res["double_free_noop"] = True
except Exception:
    res["double_free_noop"] = False
res["live_end"] = ocm.stats()["allocs_live"]
ocm.tini()
print("RESULT " + json.dumps(res))
"""


@pytest.mark.parametrize("path", ["fast", "daemon"])
def test_python_api_host_kinds(cluster2, path):
    out = _ok(run_py(cluster2, 0, API_SNIPPET, env_extra={"OCM_ALLOC_PATH": path}))
    res = json.loads(out.split("RESULT ", 1)[1])
    assert res["rank"] == 0 and res["n"] == 2 and res["dev"] == -1
    assert res["host_kind"] == ocm.OCM_LOCAL_HOST and not res["host_remote"] and res["host_rsz"] == -1
    assert res["host_zero"] == 0 and res["far_zero"] == 0            # zero-filled like the reference's calloc
    assert res["pair_kind"] == ocm.OCM_REMOTE_RDMA and res["pair_remote"]  # fixes reference ocm_is_remote bug
    assert res["pair_rank"] == 1 and res["pair_rsz"] == 2 << 20      # neighbour policy (r+1)%N
    assert res["far_eq"] and res["win_eq"] and res["off_eq"] and res["read_eq"]
    assert res["hop1"] and res["hop2"] and res["p2p_eq"]
    assert res["sig"] == 1 and res["sig2"] == 2
    assert res["self_rank"] == 0
    assert res["range_err"]
    assert res["live"] == 4 and res["live_end"] == 0
    assert (res["fast"] > 0) == (path == "fast") and (res["slow"] > 0) == (path == "daemon")
    assert res["remote_pool_used"] >= 2 << 20
    assert res["double_free_noop"]


COPY_MODEL_SNIPPET = r"""
import json, random, torch
import oncilla_b200 as ocm
ocm.init()
rng = random.Random(1234)
L, R = 6000, 9000                                  # window / far-side sizes of the pairs
loc = [ocm.alloc(ocm.OCM_LOCAL_HOST, 7000) for _ in range(2)]
pair = [ocm.alloc(ocm.OCM_REMOTE_HOST, L, R) for _ in range(2)]
# python model of every side
m_loc = [bytearray(7000) for _ in range(2)]
m_win = [bytearray(L) for _ in range(2)]
m_far = [bytearray(R) for _ in range(2)]
def fill(t, model):
    data = bytes(rng.randrange(256) for _ in range(len(model)))
    model[:] = data
    t[:] = torch.frombuffer(bytearray(data), dtype=torch.uint8)
for i in range(2):
    fill(loc[i].local_tensor(), m_loc[i]); fill(pair[i].local_tensor(), m_win[i]); fill(pair[i].remote_tensor(), m_far[i])
bad = 0; ops = 0; rejected = 0
for it in range(400):
    n = rng.randrange(1, 3000)
    so, do, so2, do2 = (rng.randrange(0, 5500) for _ in range(4))
    if rng.random() < 0.5: so2 = rng.choice([so, do])   # second hop starts where the first landed -> fused path
    kind = rng.choice(["l2l", "l2p", "p2l", "p2p", "put", "get"])
    a, b = rng.randrange(2), rng.randrange(2)
    flag = rng.choice([0, 1])
    try:
        if kind in ("put", "get"):
            op = 1 if kind == "put" else 0
            ok = n <= L and so + n <= L and do + n <= R
            (pair[a].put if op else pair[a].get)(n, src_offset=so, dest_offset=do)
            if not ok: bad += 1; continue
            if op: m_far[a][do:do+n] = m_win[a][so:so+n]
            else: m_win[a][so:so+n] = m_far[a][do:do+n]
        else:
            dst, src = {"l2l": (loc[a], loc[1-a]), "l2p": (pair[a], loc[b]), "p2l": (loc[a], pair[b]), "p2p": (pair[a], pair[1-a])}[kind]
            # what the call means once op_flag=0 has swapped the roles
            eff = kind if flag else {"l2l": "l2l", "l2p": "p2l", "p2l": "l2p", "p2p": "p2p"}[kind]
            D, S = (a, 1-a) if kind == "l2l" else (a, b) if kind in ("l2p", "p2l") else (a, 1-a)
            if not flag: D, S = S, D
            if eff == "l2l": ok = so + n <= 7000 and do + n <= 7000
            elif eff == "l2p":
                ok = so + n <= 7000 and do + n <= L and so2 + n <= L and do2 + n <= R
            elif eff == "p2l":
                ok = do + n <= R and so + n <= L and so2 + n <= L and do2 + n <= 7000
            else: ok = so + n <= R and do + n <= R
            ocm.copy(dst, src, n, src_offset=so, dest_offset=do, src_offset_2=so2, dest_offset_2=do2, op_flag=flag)
            if not ok: bad += 1; continue
            if eff == "l2l": m_loc[D][do:do+n] = m_loc[S][so:so+n]
            elif eff == "l2p":
                m_win[D][do:do+n] = m_loc[S][so:so+n]; m_far[D][do2:do2+n] = m_win[D][so2:so2+n]
            elif eff == "p2l":
                m_win[S][so:so+n] = m_far[S][do:do+n]; m_loc[D][do2:do2+n] = m_win[S][so2:so2+n]
            else: m_far[D][do:do+n] = m_far[S][so:so+n]
        ops += 1
    except ocm.OcmError as e:
        rejected += 1
        if ok: bad += 1                               # a valid request must not be rejected
def same(t, model): return bytes(t.numpy().tobytes()) == bytes(model)
for i in range(2):
    if not (same(loc[i].local_tensor(), m_loc[i]) and same(pair[i].local_tensor(), m_win[i]) and same(pair[i].remote_tensor(), m_far[i])):
        bad += 1000
for x in loc + pair: x.free()
ocm.tini()
print("RESULT " + json.dumps({"bad": bad, "ops": ops, "rejected": rejected}))
"""


def test_copy_semantics_against_model(cluster2):
    """400 random ocm_copy / put / get calls (all offset fields, both op_flags, fused and split hops, out-of-range
    requests) on host kinds, checked byte for byte against a Python model of DESIGN.md section 4."""
    out = _ok(run_py(cluster2, 0, COPY_MODEL_SNIPPET))
    res = json.loads(out.split("RESULT ", 1)[1])
    assert res["bad"] == 0, res
    assert res["ops"] > 100 and res["rejected"] > 20


SHARE_OWNER = r"""
import json, os, time, torch
import oncilla_b200 as ocm
d = os.environ["SHARE_DIR"]
ocm.init()
a = ocm.alloc(ocm.OCM_LOCAL_HOST, 1 << 20)
open(d + "/token.tmp", "wb").write(a.share()); os.rename(d + "/token.tmp", d + "/token")
a.wait_signal(2, timeout_ms=30000)                    # two puts from the peer application
t = a.local_tensor()
ok = bool((t[: 1 << 19] == 0x5A).all()) and bool((t[1 << 19 :] == 0xA5).all())
attached = a.attach_count
a.free()
open(d + "/freed", "w").write("1")
while not os.path.exists(d + "/peer_done"): time.sleep(0.01)
s = ocm.stats()
ocm.tini()
print("RESULT " + json.dumps({"ok": ok, "attached": attached, "live": s["allocs_live"]}))
"""

SHARE_PEER = r"""
import json, os, time, torch
import oncilla_b200 as ocm
d = os.environ["SHARE_DIR"]
ocm.init()
while not os.path.exists(d + "/token"): time.sleep(0.01)
att = ocm.attach(open(d + "/token", "rb").read())
info = {"kind": att.kind, "remote_bytes": att.remote_bytes, "remote_rank": att.remote_rank, "is_remote": att.is_remote}
src = ocm.alloc(ocm.OCM_LOCAL_HOST, 1 << 20)
t = src.local_tensor(); t[: 1 << 19] = 0x5A; t[1 << 19 :] = 0xA5
ocm.copy(att, src, 1 << 19)
ocm.copy(att, src, 1 << 19, src_offset=1 << 19, dest_offset=1 << 19)
info["signal"] = att.signal()
try:
    ocm.copy(att, src, (1 << 20) + 1)
    info["range"] = "accepted"
except ocm.OcmError as e:
    info["range"] = e.err
while not os.path.exists(d + "/freed"): time.sleep(0.01)
try:
    ocm.copy(att, src, 4096)
    info["after_free"] = "accepted"
except ocm.OcmError as e:
    info["after_free"] = e.err
try:
    ocm.attach(open(d + "/token", "rb").read())
    info["reattach"] = "accepted"
except ocm.OcmError as e:
    info["reattach"] = e.err
att.free(); src.free()
open(d + "/peer_done", "w").write("1")
ocm.tini()
print("RESULT " + json.dumps(info))
"""


def test_share_and_attach_between_applications(cluster2, tmp_path):
    """An application on rank 1 opens a block owned by an application on rank 0, writes it with two copies,
    and the owner's signal wait sees them; after the owner's free the attached handle is stale."""
    env = cluster2.env(0)
    env["PYTHONPATH"] = str(REPO) + os.pathsep + env.get("PYTHONPATH", "")
    env["SHARE_DIR"] = str(tmp_path)
    owner = subprocess.Popen([sys.executable, "-c", SHARE_OWNER], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                             text=True)
    try:
        peer = _ok(run_py(cluster2, 1, SHARE_PEER, env_extra={"SHARE_DIR": str(tmp_path)}))
        out, err = owner.communicate(timeout=60)
    finally:
        if owner.poll() is None:
            owner.kill()
    assert owner.returncode == 0, err
    o = json.loads(out.split("RESULT ", 1)[1])
    p = json.loads(peer.split("RESULT ", 1)[1])
    assert o == {"ok": True, "attached": 1, "live": 0}
    assert p["kind"] == ocm.OCM_REMOTE_HOST and p["remote_bytes"] == 1 << 20 and p["remote_rank"] == 0 and p["is_remote"]
    assert p["signal"] == 2
    assert p["range"] == 5                                  # OCM_ERANGE
    assert p["after_free"] == 8 and p["reattach"] == 8      # OCM_ESTALE


OVERLAP_SNIPPET = r"""
import json, random, oncilla_b200 as ocm
ocm.init()
rng = random.Random(99)
live, bad, peak, failed = [], 0, 0, 0
def overlaps(a, b):
    return a[0] == b[0] and a[1] < b[1] + b[2] and b[1] < a[1] + a[2]
for step in range(1500):
    if live and (rng.random() < 0.45 or len(live) > 60):
        h, span = live.pop(rng.randrange(len(live)))
        h.free()
        continue
    n = rng.choice([1, 100, 4096, 4097, 65536, 300000, 1 << 20])
    flags = ocm.ALLOC_NO_WINDOW | (ocm.ALLOC_SLOW_PATH if rng.random() < 0.1 else 0)
    pol = rng.choice([ocm.POLICY_DEFAULT, ocm.POLICY_SELF, ocm.POLICY_LEAST_LOADED, ocm.POLICY_ROUND_ROBIN])
    try:
        h = ocm.alloc(ocm.OCM_REMOTE_HOST, 0, n, flags=flags, policy=pol)
    except ocm.OcmError:
        failed += 1
        continue
    span = (h.remote_rank, h.remote_ptr, n)
    bad += sum(overlaps(span, s) for _, s in live)
    t = h.remote_tensor()
    bad += int(t.sum()) != 0                      # handed out zeroed, whatever lived there before
    t[:] = 0xEE
    live.append((h, span))
    peak = max(peak, len(live))
for h, _ in live: h.free()
import time; time.sleep(0.3)
left = [ocm.stats(rank=r)["pool_host_used"] for r in range(2)]
ocm.tini()
print("RESULT " + json.dumps({"bad": bad, "peak": peak, "failed": failed, "left": left}))
"""


def test_random_alloc_free_never_overlaps_and_is_zeroed(cluster2):
    """1500 random allocations and frees (all placement policies, both paths, sizes from 1 byte to 1 MiB): no two
    live blocks overlap, every block is handed out zeroed although each is dirtied before its free, and the pools
    are empty afterwards."""
    out = _ok(run_py(cluster2, 0, OVERLAP_SNIPPET))
    res = json.loads(out.split("RESULT ", 1)[1])
    assert res["bad"] == 0 and res["peak"] > 20 and res["left"] == [0, 0], res


def test_three_ranks_slow_path_forwarding():
    """REQ_ALLOC from rank 1 -> placement on rank 0 -> DO_ALLOC on rank 2 -> RELEASE_APP."""
    with ocm.LocalCluster(3, host_bytes=16 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:
        code = r"""
import json, oncilla_b200 as ocm
ocm.init()
a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20)
b = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20, policy=ocm.POLICY_EXPLICIT, rank=0)
c = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20, policy=ocm.POLICY_LEAST_LOADED)
a.local_tensor()[:] = 7
a.put(4096)
ok = int(a.remote_tensor()[:4096].sum()) == 7 * 4096
used = [ocm.stats(rank=r)["pool_host_used"] for r in range(3)]
ranks = [a.remote_rank, b.remote_rank, c.remote_rank]
for x in (a, b, c): x.free()
import time; time.sleep(0.2)
after = [ocm.stats(rank=r)["allocs_live"] for r in range(3)]
ocm.tini()
print("RESULT " + json.dumps({"ranks": ranks, "ok": ok, "used": used, "after": after}))
"""
        out = _ok(run_py(cl, 1, code, env_extra={"OCM_ALLOC_PATH": "daemon"}))
        res = json.loads(out.split("RESULT ", 1)[1])
        assert res["ranks"][0] == 2 and res["ranks"][1] == 0
        assert res["ok"]
        assert res["used"][2] >= 1 << 20 and res["used"][0] >= 1 << 20
        assert res["after"] == [0, 0, 0]


PLACE_SNIPPET = r"""
import json, oncilla_b200 as ocm
ocm.init()
res = {}
for path, flags in (("fast", 0), ("daemon", ocm.ALLOC_SLOW_PATH)):
    try:
        a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20, flags=flags)
        a.local_tensor()[:] = 5; a.put(); a.local_tensor()[:] = 0; a.get()
        res[path] = [a.remote_rank, int(a.local_tensor().sum()) == 5 * 4096]
        a.free()
    except ocm.OcmError as e:
        res[path] = ["error", e.err]
    try:
        b = ocm.alloc(ocm.OCM_REMOTE_HOST, 0, 4096, policy=ocm.POLICY_EXPLICIT, rank=1, flags=flags | ocm.ALLOC_NO_WINDOW)
        res[path + "_explicit1"] = b.remote_rank
        b.free()
    except ocm.OcmError as e:
        res[path + "_explicit1"] = ["error", e.err]
ocm.tini()
print("RESULT " + json.dumps(res))
"""


def test_dead_rank_is_skipped_and_rejoins():
    """Elastic placement: when a rank's daemon dies, its heartbeat in the cluster table goes stale and both
    allocation paths move on to the next neighbour (explicit requests for it fail with OCM_ENODAEMON instead
    of hanging like the reference); a restarted daemon re-joins and is used again."""
    with ocm.LocalCluster(3, host_bytes=16 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:
        def place():
            out = _ok(run_py(cl, 0, PLACE_SNIPPET))
            return json.loads(out.split("RESULT ", 1)[1])

        res = place()
        assert res["fast"] == [1, True] and res["daemon"] == [1, True]
        assert res["fast_explicit1"] == 1 and res["daemon_explicit1"] == 1
        cl.kill(1)
        time.sleep(2.0)                                  # heartbeat timeout is 1.5 s
        res = place()
        assert res["fast"] == [2, True] and res["daemon"] == [2, True], res
        assert res["fast_explicit1"] == ["error", 3] and res["daemon_explicit1"] == ["error", 3], res   # OCM_ENODAEMON
        cl.restart(1)
        time.sleep(0.3)
        res = place()
        assert res["fast"] == [1, True] and res["daemon"] == [1, True], res
        assert res["fast_explicit1"] == 1 and res["daemon_explicit1"] == 1


SURVIVOR = r"""
import json, os, time, oncilla_b200 as ocm
d = os.environ["SYNC_DIR"]
ocm.init()
a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20)              # far side on rank 1, first incarnation
a.local_tensor()[:] = 9; a.put()
res = {"a_rank": a.remote_rank}
open(d + "/allocated", "w").write("1")
while not os.path.exists(d + "/restarted"): time.sleep(0.01)
b = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20)              # must land in the NEW pool of rank 1
res["b_rank"] = b.remote_rank
b.local_tensor()[:] = 4; b.put(); b.local_tensor()[:] = 0; b.get()
res["b_ok"] = int(b.local_tensor().sum()) == 4 * 4096
res["b_zero_tail"] = int(b.remote_tensor()[4096:].sum()) == 0   # fresh, zeroed memory (not a's 9s)
try:
    a.free(); res["a_free"] = "ok"
except ocm.OcmError as e:
    res["a_free"] = e.err
res["used_mid"] = ocm.stats(rank=1)["pool_host_used"]
b.free()
time.sleep(0.2)
res["live_end"] = ocm.stats(rank=1)["allocs_live"]
ocm.tini()
print("RESULT " + json.dumps(res))
"""


def test_application_survives_peer_daemon_restart(tmp_path):
    """A long-lived application keeps working when the daemon of the rank that backs its memory is restarted:
    new allocations land in the new incarnation's pool, handles from before the restart go stale instead of
    corrupting the new tables, and the capacity accounting starts from zero."""
    with ocm.LocalCluster(2, host_bytes=16 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:
        env = cl.env(0)
        env["PYTHONPATH"] = str(REPO) + os.pathsep + env.get("PYTHONPATH", "")
        env["SYNC_DIR"] = str(tmp_path)
        app = subprocess.Popen([sys.executable, "-c", SURVIVOR], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                               text=True)
        try:
            t0 = time.time()
            while not (tmp_path / "allocated").exists():
                assert app.poll() is None and time.time() - t0 < 60, app.stderr.read() if app.poll() is not None else "timeout"
                time.sleep(0.02)
            cl.restart(1)
            time.sleep(0.3)
            (tmp_path / "restarted").write_text("1")
            out, err = app.communicate(timeout=60)
        finally:
            if app.poll() is None:
                app.kill()
        assert app.returncode == 0, err
        res = json.loads(out.split("RESULT ", 1)[1])
        assert res["a_rank"] == 1 and res["b_rank"] == 1 and res["b_ok"] and res["b_zero_tail"]
        assert res["a_free"] == 8                                  # OCM_ESTALE: the block died with the old daemon
        assert res["used_mid"] == 1 << 20                           # only b counts in the new incarnation
        assert res["live_end"] == 0


CKPT_SAVE = r"""
import json, os, torch, oncilla_b200 as ocm
from oncilla_b200 import checkpoint as ckpt
path = os.environ["CKPT"]
ocm.init()
g = torch.Generator().manual_seed(11)
a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, (3 << 20) + 123)          # payload = far side on rank 1
b = ocm.alloc(ocm.OCM_LOCAL_HOST, 70000)                            # payload = local memory
a.remote_tensor().copy_(torch.randint(0, 256, (a.remote_bytes,), dtype=torch.uint8, generator=g))
b.local_tensor().copy_(torch.randint(0, 256, (70000,), dtype=torch.uint8, generator=g))
hdr = ckpt.save(path, {"far": a, "near": b})
print("RESULT " + json.dumps({"entries": [[e["name"], e["bytes"]] for e in hdr["entries"]],
                              "sum": [int(a.remote_tensor().sum()), int(b.local_tensor().sum())]}))
os._exit(0)                                                         # the application goes away without cleaning up
"""

CKPT_LOAD = r"""
import json, os, torch, oncilla_b200 as ocm
from oncilla_b200 import checkpoint as ckpt
path = os.environ["CKPT"]
ocm.init()
a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, (3 << 20) + 123)
b = ocm.alloc(ocm.OCM_LOCAL_HOST, 70000)
res = {"zero_before": int(a.remote_tensor().sum()) == 0 and int(b.local_tensor().sum()) == 0}
ckpt.load(path, {"far": a, "near": b})
res["sum"] = [int(a.remote_tensor().sum()), int(b.local_tensor().sum())]
res["names"] = [e["name"] for e in ckpt.read_header(path)["entries"]]
errs = []
small = ocm.alloc(ocm.OCM_LOCAL_HOST, 100)
for what, fn in (("size", lambda: ckpt.load(path, {"near": small})), ("name", lambda: ckpt.load(path, {"nope": b}))):
    try: fn(); errs.append(what + ":accepted")
    except ckpt.CheckpointError: errs.append(what + ":rejected")
raw = bytearray(open(path, "rb").read()); raw[(1 << 20) + 5000] ^= 0xFF; open(path + ".bad", "wb").write(raw)
try: ckpt.load(path + ".bad", {"far": a}); errs.append("crc:accepted")
except ckpt.CheckpointError: errs.append("crc:rejected")
res["errs"] = errs
for x in (a, b, small): x.free()
ocm.tini()
print("RESULT " + json.dumps(res))
"""


def test_checkpoint_survives_cluster_restart(tmp_path):
    """oncilla_b200.checkpoint: payloads of a pair's far side and of a local buffer are saved, the whole cluster is
    torn down, a new one restores them into fresh allocations; size, name and CRC mismatches are rejected."""
    path = str(tmp_path / "state.ocm")
    with ocm.LocalCluster(2, host_bytes=32 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:
        out = _ok(run_py(cl, 0, CKPT_SAVE, env_extra={"CKPT": path}))
        saved = json.loads(out.split("RESULT ", 1)[1])
    assert saved["entries"] == [["far", (3 << 20) + 123], ["near", 70000]]
    with ocm.LocalCluster(2, host_bytes=32 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:       # brand-new daemons
        out = _ok(run_py(cl, 0, CKPT_LOAD, env_extra={"CKPT": path}))
        res = json.loads(out.split("RESULT ", 1)[1])
    assert res["zero_before"] and res["sum"] == saved["sum"] and res["names"] == ["far", "near"]
    assert res["errs"] == ["size:rejected", "name:rejected", "crc:rejected"]


def test_torchrun_cluster_all_to_all_over_shared_blocks():
    """Three processes under torchrun (gloo, no GPU): TorchrunCluster brings up one daemon per rank, every rank shares
    a receive block, attaches to the others' and writes its slot everywhere; signals count every sender."""
    env = {k: v for k, v in os.environ.items() if not k.startswith("OCM_")}
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3",
                        "--master-addr", "127.0.0.1", "--master-port", "29933", str(REPO / "tests" / "torchrun_exchange.py")],
                       env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    res = json.loads(r.stdout.split("RESULT ", 1)[1].splitlines()[0])
    assert res == [[True, 1, 3], [True, 2, 3], [True, 0, 3]], res


def test_duplicate_daemon_cannot_hurt_a_live_cluster(cluster2):
    """Starting a second daemon for a rank that is being served fails at once - before it replaces the live
    daemon's tables, host pool or message queue."""
    from oncilla_b200 import _native

    for rank in (0, 1):
        r = subprocess.run([str(_native.DAEMON_PATH), cluster2.nodefile], env=cluster2.env(rank), capture_output=True,
                           text=True, timeout=30)
        assert r.returncode != 0 and "already served by a running daemon" in (r.stdout + r.stderr)
    out = _ok(run_py(cluster2, 1, PLACE_SNIPPET))
    res = json.loads(out.split("RESULT ", 1)[1])
    assert res["fast"] == [0, True] and res["daemon"] == [0, True], res


def test_cleanup_removes_only_what_dead_daemons_left(cluster2):
    """python -m oncilla_b200.cleanup: segments and queues of SIGKILLed daemons go, a live cluster stays intact."""
    from oncilla_b200 import cleanup

    doomed = ocm.LocalCluster(2, host_bytes=8 << 20, extra_env={"OCM_NO_GPU": "1"}, ns=f"doomed{os.getpid()}")
    doomed.start()
    ns = doomed.ns
    try:
        assert cleanup.sweep(ns) == []                                  # alive: nothing to do
        for r in (1, 0):
            doomed.kill(r)                                              # SIGKILL: no chance to unlink anything
        left = sorted(p.name for p in cleanup.SHM.iterdir() if p.name.startswith(f"ocm_{ns}_"))
        assert f"ocm_{ns}_tbl_r0" in left and f"ocm_{ns}_host_r1" in left and f"ocm_{ns}_cluster" in left
        assert set(cleanup.sweep(ns, dry_run=True)) >= set(left)
        removed = cleanup.sweep(ns)
        assert set(removed) >= set(left) and f"/ocm_mq_{ns}_daemon" in removed
        assert not [p for p in cleanup.SHM.iterdir() if p.name.startswith(f"ocm_{ns}_")]
    finally:
        doomed.stop()
    assert cleanup.sweep(cluster2.ns) == []                             # the fixture's live cluster is untouched
    out = _ok(run_py(cluster2, 0, "import oncilla_b200 as ocm; ocm.init(); a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 4096); a.put(); a.free(); ocm.tini(); print('STILL_OK')"))
    assert "STILL_OK" in out


def test_app_death_is_reclaimed(cluster2):
    """A crashed application's regions return to the pool (TODO in reference README:68-69)."""
    code = r"""
import os, oncilla_b200 as ocm
ocm.init()
keep = [ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 4 << 20) for _ in range(3)]
print("USED", ocm.stats(rank=1)["pool_host_used"], flush=True)
os._exit(0)   # no free, no tini
"""
    out = _ok(run_py(cluster2, 0, code))
    assert int(out.split("USED ", 1)[1].split()[0]) >= 12 << 20
    probe = r"""
import time, oncilla_b200 as ocm
ocm.init()
t0 = time.time()
while time.time() - t0 < 10:
    s0, s1 = ocm.stats(rank=0), ocm.stats(rank=1)
    if s0["allocs_live"] == 0 and s1["allocs_live"] == 0 and s1["pool_host_used"] == 0: break
    time.sleep(0.05)
print("LIVE", s0["allocs_live"], s1["allocs_live"], s1["pool_host_used"])
a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 24 << 20)   # needs the reclaimed space (32 MiB pool)
print("ZERO", int(a.remote_tensor().sum()))
a.free(); ocm.tini()
"""
    out = _ok(run_py(cluster2, 0, probe))
    assert "LIVE 0 0 0" in out and "ZERO 0" in out


def test_daemon_death_is_detected(cluster2):
    """Calls fail promptly when a daemon disappears instead of spinning forever (reference src/pmsg.c:134-151)."""
    code = r"""
import os, signal, time, json, oncilla_b200 as ocm
ocm.init()
a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 1 << 20)
os.kill(int(os.environ["VICTIM"]), signal.SIGKILL)
time.sleep(0.3)
t0 = time.time()
try:
    ocm.stats(rank=0); ok = False
except ocm.OcmError as e:
    ok = e.err == 3
print("RESULT " + json.dumps({"ok": ok, "dt": time.time() - t0}))
os._exit(0)
"""
    victim = cluster2.procs[0].pid
    out = _ok(run_py(cluster2, 0, code, env_extra={"VICTIM": str(victim)}))
    res = json.loads(out.split("RESULT ", 1)[1])
    assert res["ok"] and res["dt"] < 5


def test_exhaustion_and_local_fallback():
    with ocm.LocalCluster(1, host_bytes=4 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:
        code = r"""
import json, oncilla_b200 as ocm
ocm.init()
res = {}
try:
    ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 64 << 20); res["enomem"] = False
except ocm.OcmError as e:
    res["enomem"] = e.err == 2
big = ocm.alloc(ocm.OCM_LOCAL_HOST, 16 << 20)      # larger than the pool: malloc fallback like the reference
big.local_tensor()[:] = 1
res["fallback_ok"] = int(big.local_tensor().sum()) == 16 << 20
one = ocm.alloc(ocm.OCM_REMOTE_RMA, 4096, 1 << 20)  # single-rank cluster: stays on rank 0
res["single_rank"] = one.remote_rank
big.free(); one.free(); ocm.tini()
print("RESULT " + json.dumps(res))
"""
        out = _ok(run_py(cl, 0, code))
        res = json.loads(out.split("RESULT ", 1)[1])
        assert res == {"enomem": True, "fallback_ok": True, "single_rank": 0}


def test_many_clients_alloc_free(cluster2):
    """Scaled-down many-client stress (BASELINE.json config 5: 64 apps x 10k; the full size runs in bench/)."""
    n_apps, iters = 12, 400
    procs = []
    for i in range(n_apps):
        env = cluster2.env(i % 2)
        env["OCM_ALLOC_PATH"] = "fast" if i % 3 else "daemon"
        procs.append(subprocess.Popen([str(BIN / "ocm_stress"), str(iters), "65536"], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "pass" in o
    code = r"""
import oncilla_b200 as ocm, time
ocm.init(); time.sleep(0.3)
print("LIVE", [ocm.stats(rank=r)["allocs_live"] for r in range(2)], [ocm.stats(rank=r)["pool_host_used"] for r in range(2)])
ocm.tini()
"""
    out = _ok(run_py(cluster2, 0, code))
    assert "LIVE [0, 0] [0, 0]" in out


def test_fault_injection_keeps_accounting_consistent():
    """Every 3rd owner-side fulfilment fails (OCM_FAULT_ALLOC_EVERY): errors surface, reservations are returned."""
    with ocm.LocalCluster(2, host_bytes=16 << 20, extra_env={"OCM_NO_GPU": "1", "OCM_FAULT_ALLOC_EVERY": "3"}) as cl:
        code = r"""
import json, time, oncilla_b200 as ocm
ocm.init()
ok = fail = 0
held = []
for i in range(30):
    try:
        held.append(ocm.alloc(ocm.OCM_REMOTE_HOST, 0, 64 << 10, flags=ocm.ALLOC_NO_WINDOW | ocm.ALLOC_SLOW_PATH)); ok += 1
    except ocm.OcmError as e:
        assert e.err == 2, e
        fail += 1
for a in held: a.free()
time.sleep(0.3)
live = [ocm.stats(rank=r)["allocs_live"] for r in range(2)]
used = [ocm.stats(rank=r)["pool_host_used"] for r in range(2)]
big = ocm.alloc(ocm.OCM_REMOTE_HOST, 0, 12 << 20, flags=ocm.ALLOC_NO_WINDOW)   # capacity accounting did not leak
big.free(); ocm.tini()
print("RESULT " + json.dumps({"ok": ok, "fail": fail, "live": live, "used": used}))
"""
        out = _ok(run_py(cl, 0, code))
        res = json.loads(out.split("RESULT ", 1)[1])
        assert res["ok"] == 20 and res["fail"] == 10
        assert res["live"] == [0, 0] and res["used"] == [0, 0]


def test_reference_format_nodefile(tmp_path):
    """A nodefile in the reference's exact 5-column format boots a daemon when OCM_RANK selects the line."""
    nf = tmp_path / "nodefile"
    ports = ocm.cluster._free_ports(1)
    nf.write_text(f"#rank dns ethernet_ip ocm_port rdmacm_port\n0 shiva.cc.gt.atl.ga.us 127.0.0.1 {ports[0]} 67890\n")
    with ocm.LocalCluster(1, nodefile=str(nf), extra_env={"OCM_NO_GPU": "1", "OCM_HOST_BYTES": "8M"}) as cl:
        out = _ok(run_app(cl, 0, [BIN / "ocm_test", 1, "1", "1", "1"]))
        assert "pass" in out


@pytest.mark.parametrize("node_local", ["1", "0"])
def test_daemon_path_node_local_and_message_path(node_local):
    """Slow-path allocations from rank 1's application: with OCM_DAEMON_NODE_LOCAL=1 its daemon places and fulfils them
    in the shared tables of this node, with =0 the request travels rank 1 -> rank 0 (placement) -> owner (fulfilment)
    over TCP as between nodes (reference src/main.c + src/alloc.c flow).  Same results, same accounting."""
    env = {"OCM_NO_GPU": "1", "OCM_DAEMON_NODE_LOCAL": node_local}
    with ocm.LocalCluster(3, host_bytes=16 << 20, extra_env=env) as cl:
        code = r"""
import json, time, oncilla_b200 as ocm
ocm.init()
held, ranks = [], set()
for i in range(40):
    a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 100000 + i, flags=ocm.ALLOC_SLOW_PATH)
    ranks.add(a.remote_rank)
    a.local_tensor()[:16] = i
    a.put(16)
    held.append(a)
mid = [ocm.stats(rank=r)["allocs_live"] for r in range(3)]
for i, a in enumerate(held):
    a.local_tensor()[:16] = 255
    a.get(16)
    assert int(a.local_tensor()[3]) == i
    a.free()
time.sleep(0.3)
end = [ocm.stats(rank=r)["allocs_live"] for r in range(3)]
used = [ocm.stats(rank=r)["pool_host_used"] for r in range(3)]
me = ocm.stats()
ocm.tini()
print("RESULT " + json.dumps({"ranks": sorted(ranks), "mid": mid, "end": end, "used": used, "slow": me.get("allocs_slow", -1)}))
"""
        out = _ok(run_py(cl, 1, code))
        res = json.loads(out.split("RESULT ", 1)[1])
        assert res["ranks"] == [2], res                      # neighbour policy from rank 1
        assert sum(res["mid"]) >= 40 and res["end"] == [0, 0, 0] and res["used"] == [0, 0, 0], res


def test_prefetch_views_maps_every_rank(cluster2):
    """ocm_prefetch_views maps all ranks' tables and host pools at once (set-up time) instead of at first use."""
    code = r"""
import oncilla_b200 as ocm
ocm.init()
n = ocm.prefetch_views(gpu=False, host=True)
assert n == ocm.num_ranks() == 2, n
assert ocm.prefetch_views(gpu=False, host=True) == 2          # idempotent
a = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 4096); a.put(64); a.free()
ocm.tini()
print("PREFETCH OK")
"""
    assert "PREFETCH OK" in _ok(run_py(cluster2, 0, code))
