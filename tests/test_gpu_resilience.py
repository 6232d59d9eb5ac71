"""GPU-side versions of the resilience tests: checkpoint / restore and daemon-restart survival with
payloads in HBM, the TMA bulk engine on peer memory, and one transfer beyond 4 GiB (the reference's raw bandwidth
test allocates 4 GiB + 1 and its sge.length is 32-bit: reference test/ib_client.c:85-108, src/rdma.c:61-63).
"""
import json
import os
import subprocess
import sys
import time

import pytest
import torch

import oncilla_b200 as ocm
from conftest import REPO, run_py

pytestmark = pytest.mark.gpu

N_GPUS = min(torch.cuda.device_count(), 2) if torch.cuda.is_available() else 0


def _cluster(**kw):
    return ocm.LocalCluster(N_GPUS, gpus=list(range(N_GPUS)), **kw)


def _ok(r):
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    return r.stdout


CKPT_SAVE = r"""
import json, os, torch, oncilla_b200 as ocm
from oncilla_b200 import checkpoint as ckpt
path = os.environ["CKPT"]
ocm.init()
g = torch.Generator().manual_seed(11)
a = ocm.alloc(ocm.OCM_REMOTE_GPU, 4096, (3 << 20) + 123)           # payload = far side in the neighbour's HBM pool
b = ocm.alloc(ocm.OCM_LOCAL_GPU, 70000)                             # payload = our own HBM pool
a.remote_tensor().copy_(torch.randint(0, 256, (a.remote_bytes,), dtype=torch.uint8, generator=g).cuda())
b.local_tensor().copy_(torch.randint(0, 256, (70000,), dtype=torch.uint8, generator=g).cuda())
torch.cuda.synchronize()
hdr = ckpt.save(path, {"far": a, "near": b})
print("RESULT " + json.dumps({"entries": [[e["name"], e["bytes"]] for e in hdr["entries"]], "rank": a.remote_rank,
                              "sum": [int(a.remote_tensor().sum()), int(b.local_tensor().sum())]}))
os._exit(0)                                                         # the application goes away without cleaning up
"""

CKPT_LOAD = r"""
import json, os, torch, oncilla_b200 as ocm
from oncilla_b200 import checkpoint as ckpt
path = os.environ["CKPT"]
ocm.init()
a = ocm.alloc(ocm.OCM_REMOTE_GPU, 4096, (3 << 20) + 123)
b = ocm.alloc(ocm.OCM_LOCAL_GPU, 70000)
res = {"zero_before": int(a.remote_tensor().sum()) == 0 and int(b.local_tensor().sum()) == 0, "rank": a.remote_rank}
ckpt.load(path, {"far": a, "near": b})
res["sum"] = [int(a.remote_tensor().sum()), int(b.local_tensor().sum())]
raw = bytearray(open(path, "rb").read()); raw[(1 << 20) + 5000] ^= 0xFF; open(path + ".bad", "wb").write(raw)
try: ckpt.load(path + ".bad", {"far": a}); res["crc"] = "accepted"
except ckpt.CheckpointError: res["crc"] = "rejected"
for x in (a, b): x.free()
ocm.tini()
print("RESULT " + json.dumps(res))
"""


def test_checkpoint_restores_hbm_payloads_into_a_new_cluster(tmp_path):
    """Payloads that live in HBM pools (a pair's far side on the neighbour GPU, a local GPU buffer) are saved, every
    daemon is torn down - the pools die with them - and a brand-new cluster restores them byte for byte."""
    path = str(tmp_path / "state.ocm")
    with _cluster(gpu_bytes=256 << 20, host_bytes=32 << 20) as cl:
        saved = json.loads(_ok(run_py(cl, 0, CKPT_SAVE, env_extra={"CKPT": path})).split("RESULT ", 1)[1])
    assert saved["entries"] == [["far", (3 << 20) + 123], ["near", 70000]] and saved["rank"] == (1 % N_GPUS)
    with _cluster(gpu_bytes=256 << 20, host_bytes=32 << 20) as cl:
        res = json.loads(_ok(run_py(cl, 0, CKPT_LOAD, env_extra={"CKPT": path})).split("RESULT ", 1)[1])
    assert res["zero_before"] and res["sum"] == saved["sum"] and res["crc"] == "rejected" and res["rank"] == saved["rank"]


SURVIVOR = r"""
import json, os, time, torch, oncilla_b200 as ocm
from oncilla_b200 import ops
d = os.environ["SYNC_DIR"]
ocm.init()
n = 1 << 20
a = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)                            # far side in rank 1's HBM, first incarnation
ops.fill(a.local_tensor(), 0x0909090909090909); torch.cuda.synchronize(); a.put()
res = {"a_rank": a.remote_rank}
open(d + "/allocated", "w").write("1")
while not os.path.exists(d + "/restarted"): time.sleep(0.01)
b = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)                            # must land in the NEW pool of rank 1
res["b_rank"] = b.remote_rank
res["b_zero"] = int(b.remote_tensor().sum()) == 0                  # fresh, zeroed HBM (not a's 9s)
ops.fill(b.local_tensor(), 0x0404040404040404); torch.cuda.synchronize()
b.put(); ops.fill(b.local_tensor(), 0); torch.cuda.synchronize(); b.get()
res["b_ok"] = ops.verify(b.local_tensor(), 0x0404040404040404) == 0
try:
    a.free(); res["a_free"] = "ok"
except ocm.OcmError as e:
    res["a_free"] = e.err
b.free()
time.sleep(0.3)
res["live_end"] = ocm.stats(rank=1)["allocs_live"]
ocm.tini()
print("RESULT " + json.dumps(res))
"""


@pytest.mark.skipif(N_GPUS < 2, reason="needs two GPUs (the far side must live in another rank's pool)")
def test_application_survives_restart_of_the_rank_that_holds_its_hbm(tmp_path):
    """The daemon that owns the HBM pool behind a live pair is killed and restarted: the application maps the new
    incarnation's pool for new allocations, the old handle goes stale (OCM_ESTALE) instead of touching the new tables."""
    with _cluster(gpu_bytes=256 << 20, host_bytes=32 << 20) as cl:
        env = cl.env(0)
        env["PYTHONPATH"] = str(REPO) + os.pathsep + env.get("PYTHONPATH", "")
        env["SYNC_DIR"] = str(tmp_path)
        app = subprocess.Popen([sys.executable, "-c", SURVIVOR], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        try:
            t0 = time.time()
            while not (tmp_path / "allocated").exists():
                assert app.poll() is None and time.time() - t0 < 120, app.stderr.read() if app.poll() is not None else "timeout"
                time.sleep(0.02)
            cl.restart(1)
            time.sleep(0.5)
            (tmp_path / "restarted").write_text("1")
            out, err = app.communicate(timeout=120)
        finally:
            if app.poll() is None:
                app.kill()
        assert app.returncode == 0, err[-3000:]
        res = json.loads(out.split("RESULT ", 1)[1])
        assert res["a_rank"] == 1 and res["b_rank"] == 1 and res["b_zero"] and res["b_ok"], res
        assert res["a_free"] == 8 and res["live_end"] == 0, res      # 8 = OCM_ESTALE


BULK_PEER = r"""
import torch, oncilla_b200 as ocm
from oncilla_b200 import ops
ocm.init()
for n in (4096, (1 << 20) + 16, 48 << 20):
    a = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
    assert a.remote_rank == 1
    pat = 0xB01CB01C00000000 | n
    ops.fill(a.local_tensor(), pat); torch.cuda.synchronize()
    a.put(n, stream=None); ocm.quiet()
    assert ops.verify(a.remote_tensor(), pat, n) == 0, n
    ops.fill(a.local_tensor(), 0); torch.cuda.synchronize()
    a.get(n, stream=None); ocm.quiet()
    assert ops.verify(a.local_tensor(), pat, n) == 0, n
    src = ocm.alloc(ocm.OCM_LOCAL_GPU, n)
    ops.fill(src.local_tensor(), pat ^ 0xFFFF); torch.cuda.synchronize()
    ocm.copy(a, src, n, stream=None); ocm.quiet()                 # two destinations: window + far side
    assert ops.verify(a.remote_tensor(), pat ^ 0xFFFF, n) == 0 and ops.verify(a.local_tensor(), pat ^ 0xFFFF, n) == 0
    assert a.signal() == 2
    a.free(); src.free()
ocm.tini()
print("BULK PEER OK")
"""


@pytest.mark.skipif(N_GPUS < 2, reason="needs two GPUs")
def test_tma_bulk_engine_over_nvlink():
    """OCM_COPY_MODE=bulk (cp.async.bulk staging through shared memory, SASS UBLKCP) with the far side in a PEER GPU's
    pool: put, get and the two-destination copy, with completion signals."""
    with _cluster(gpu_bytes=512 << 20, host_bytes=32 << 20) as cl:
        out = _ok(run_py(cl, 0, BULK_PEER, env_extra={"OCM_COPY_MODE": "bulk"}))
        assert "BULK PEER OK" in out


def test_one_transfer_beyond_4_gib():
    """4 GiB + 1 bytes in ONE ocm_copy_onesided each way (all offsets and lengths are 64-bit end to end; the
    reference's transport carries a 32-bit length): first, last and straddling bytes are checked."""
    n = (4 << 30) + 1
    free_b, _ = torch.cuda.mem_get_info(0)
    if free_b < 3 * n + (2 << 30):
        pytest.skip("not enough free HBM for a 4 GiB + 1 window and far side")
    with _cluster(gpu_bytes=2 * n + (512 << 20), host_bytes=32 << 20) as cl:
        code = r"""
import torch, oncilla_b200 as ocm
from oncilla_b200 import ops
ocm.init()
n = (4 << 30) + 1
a = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
w, f = a.local_tensor(), a.remote_tensor()
spots = [0, 1, (1 << 31) - 1, 1 << 31, (1 << 32) - 1, n - 1, (2 << 30) + 12345, (3 << 30) + 7]   # n - 1 == 1 << 32
ops.fill(w[: n - 1], 0x5151515151515151)
for i, s in enumerate(spots): w[s] = 200 + i
torch.cuda.synchronize()
a.put(n)
assert [int(f[s]) for s in spots] == [200 + i for i in range(len(spots))]
assert int(f[(1 << 32) - 2]) == 0x51 and int(f[(1 << 31) + 1]) == 0x51 and int(f[2]) == 0x51
assert ops.verify(f[8:1 << 30], 0x5151515151515151, (1 << 30) - 8) == 0
w.zero_(); torch.cuda.synchronize()
a.get(n)
assert [int(w[s]) for s in spots] == [200 + i for i in range(len(spots))]
assert int(w[(1 << 32) - 2]) == 0x51 and int(w[(1 << 31) + 1]) == 0x51
a.free(); ocm.tini()
print("BEYOND 4 GIB OK")
"""
        assert "BEYOND 4 GIB OK" in _ok(run_py(cl, 0, code, timeout=300))


REPL_NO_NVLS = r"""
import torch, oncilla_b200 as ocm
from oncilla_b200 import ops
ocm.init()
assert not ocm.multicast_ready(0.3)
N = ocm.num_ranks()
for n in (4096, (5 << 20) + 48):
    rp = ocm.alloc(ocm.OCM_REPLICATED_GPU, n, n)
    pat = 0x7E57000000000000 | n
    ops.fill(rp.local_tensor(), pat); torch.cuda.synchronize()
    rp.put(n)                                                       # source = our own pool: cluster-shared tiles
    for r in range(N):
        assert ops.verify(rp.slab_tensor(r)[:n], pat, n) == 0, (n, r)
    far = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)                       # far side in the neighbour's HBM
    ops.fill(far.local_tensor(), pat ^ 0xABCD); torch.cuda.synchronize(); far.put(n)
    dsts = [rp.slab_tensor(r)[:n] for r in range(N)]
    ops.put_replicated(dsts, far.remote_tensor()[:n], n)            # source = a peer's memory: per-CTA loads
    torch.cuda.synchronize()
    for r in range(N):
        assert ops.verify(rp.slab_tensor(r)[:n], pat ^ 0xABCD, n) == 0, (n, r)
    rp.free(); far.free()
ocm.tini()
print("REPLICATED WITHOUT NVLS OK")
"""


def test_replicated_put_without_nvls_local_and_peer_source():
    """OCM_MC_BYTES=0 keeps the NVLS region out, so OCM_REPLICATED_GPU puts run the cluster kernel (K6a): from this
    GPU's own pool, and - through the kernel ABI - with a source that lives in the neighbour's HBM."""
    with _cluster(gpu_bytes=256 << 20, host_bytes=32 << 20, extra_env={"OCM_MC_BYTES": "0"}) as cl:
        assert "REPLICATED WITHOUT NVLS OK" in _ok(run_py(cl, 0, REPL_NO_NVLS))


HOST_PIECES = r"""
import torch, oncilla_b200 as ocm
from oncilla_b200 import ops
ocm.init()
dev = torch.device(f"cuda:{ocm.device()}")
g = ocm.alloc(ocm.OCM_LOCAL_GPU, 9 << 20)
blocks = []
for i in range(6):                                   # 5 MiB blocks in a pool pinned in 2 MiB pieces: each spans 3-4 pieces,
    h = ocm.alloc(ocm.OCM_REMOTE_HOST, 0, 5 << 20, flags=ocm.ALLOC_NO_WINDOW)   # registered at different times
    blocks.append(h)
for i, h in enumerate(blocks):
    pat = 0xC0FFEE0000000000 | i
    ops.fill(g.local_tensor()[: 5 << 20], pat); torch.cuda.synchronize()
    ocm.copy(h, g, 5 << 20)                          # GPU -> host block (copy engine, cut at the piece boundaries)
    ops.fill(g.local_tensor(), 0); torch.cuda.synchronize()
    ocm.copy(g, h, (5 << 20) - 4096, src_offset=4096, dest_offset=8192)
    assert ops.verify(g.local_tensor()[8192: 8192 + (5 << 20) - 4096], pat, (5 << 20) - 4096) == 0, i
blocks[1].free(); blocks[3].free()
big = ocm.alloc(ocm.OCM_REMOTE_HOST, 0, 9 << 20, flags=ocm.ALLOC_NO_WINDOW)     # lands across old and new pieces
ops.fill(g.local_tensor(), 0x77); torch.cuda.synchronize()
ocm.copy(big, g, 9 << 20)
ops.fill(g.local_tensor(), 0); torch.cuda.synchronize()
ocm.copy(g, big, 9 << 20)
assert ops.verify(g.local_tensor(), 0x77, 9 << 20) == 0
for h in (blocks[0], blocks[2], blocks[4], blocks[5], big, g): h.free()
ocm.tini()
print("HOST PIECES OK")
"""


def test_host_pool_pinned_in_pieces():
    """OCM_HOST_REG_CHUNK=2M: a process pins a rank's host pool piece by piece as its blocks are used; blocks that span
    pieces registered at different times still copy to and from the GPU (the library cuts its copies at the seams)."""
    with _cluster(gpu_bytes=256 << 20, host_bytes=64 << 20, extra_env={"OCM_HOST_POOL": "shm"}) as cl:
        assert "HOST PIECES OK" in _ok(run_py(cl, 0, HOST_PIECES, env_extra={"OCM_HOST_REG_CHUNK": str(2 << 20)}))


def test_host_pool_is_a_pinned_vmm_allocation_by_default():
    """With a GPU the daemon creates its host pool as a pinned CUDA VMM allocation and hands it out by descriptor:
    the application maps it without pinning anything, CPU and GPU both see the bytes, fresh blocks are zero."""
    code = r"""
import torch, oncilla_b200 as ocm
from oncilla_b200 import ops
ocm.init()
h = ocm.alloc(ocm.OCM_REMOTE_HOST, 1 << 20, 5 << 20)
f = h.remote_tensor()
assert int(f.sum()) == 0                                    # CPU reads zeros
ops.fill(h.local_tensor(), 0x1122334455667788); h.put(1 << 20)
assert ops.verify(f[: 1 << 20], 0x1122334455667788, 1 << 20) == 0
g = ocm.alloc(ocm.OCM_LOCAL_GPU, 1 << 20)
ocm.copy(g, h, 1 << 20, src_offset=0)                      # host window -> GPU through the library
assert ops.verify(g.local_tensor(), 0x1122334455667788, 1 << 20) == 0
h.free(); g.free(); ocm.tini()
print("VMM HOST OK")
"""
    with _cluster(gpu_bytes=256 << 20, host_bytes=64 << 20, log_level="info") as cl:
        assert "VMM HOST OK" in _ok(run_py(cl, 0, code))
        assert "pinned vmm-fd" in cl.read_log(0), cl.read_log(0)[-600:]
