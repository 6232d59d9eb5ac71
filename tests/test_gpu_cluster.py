"""End-to-end on a real GPU: daemons own HBM pools (CUDA VMM, fd-exported), libocm maps them, every copy
runs through our kernels.  Mirrors the reference's ocm_test scenarios (reference test/ocm_test.c) with the
GPU kinds that the reference leaves unimplemented (OCM_REMOTE_GPU is rejected there: src/lib.c:207-211,
src/alloc.c:204-208) plus verification of every transfer.
"""
import ctypes
import subprocess

import pytest
import torch

import oncilla_b200 as ocm
from conftest import BIN, run_app, run_py
from oncilla_b200 import ops

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu_cluster():
    """As many ranks as GPUs (max 2 to keep it cheap); this process attaches to rank 0."""
    n = min(torch.cuda.device_count(), 2)
    torch.cuda.set_device(0)
    cl = ocm.LocalCluster(n, gpus=list(range(n)), gpu_bytes=1 << 30, host_bytes=256 << 20, ns=None)
    cl.start()
    cl.apply_env(0)
    ocm.init()
    yield cl
    ocm.tini()
    cl.stop()


def test_attach_and_kinds(gpu_cluster):
    assert ocm.rank() == 0 and ocm.num_ranks() == gpu_cluster.n_ranks and ocm.device() == 0
    g = ocm.alloc(ocm.OCM_LOCAL_GPU, 4 << 20)
    assert g.kind == ocm.OCM_LOCAL_GPU and not g.is_remote and g.remote_bytes == -1
    t = g.local_tensor()
    assert t.is_cuda and t.numel() == 4 << 20 and int(t.sum()) == 0     # zero on alloc
    r = ocm.alloc(ocm.OCM_REMOTE_GPU, 2 << 20, 8 << 20)
    assert r.kind == ocm.OCM_REMOTE_GPU and r.is_remote and r.remote_bytes == 8 << 20
    assert r.remote_rank == (1 if gpu_cluster.n_ranks > 1 else 0)        # neighbour policy
    assert int(r.remote_tensor().sum()) == 0
    st = ocm.stats()
    assert st["fast_path_allocs"] >= 3 and st["pool_gpu_bytes"] >= 1 << 30
    g.free()
    r.free()


@pytest.mark.parametrize("nbytes", [64, 4096, (1 << 20) + 16, 32 << 20])
def test_put_get_roundtrip(gpu_cluster, nbytes):
    a = ocm.alloc(ocm.OCM_REMOTE_GPU, nbytes, nbytes)
    w = a.local_tensor()
    pat = 0xDEADBEEF00000000 | nbytes
    ops.fill(w, pat)
    torch.cuda.synchronize()
    a.put(nbytes)
    assert a.signal() == 1
    assert ops.verify(a.remote_tensor(), pat, nbytes) == 0              # the bytes are in the peer pool
    ops.fill(w, 0)
    torch.cuda.synchronize()
    a.get(nbytes)
    assert ops.verify(w, pat, nbytes) == 0
    a.free()


def test_copy_matrix_with_gpu(gpu_cluster):
    n = 3 << 20
    h1, h2 = ocm.alloc(ocm.OCM_LOCAL_HOST, n), ocm.alloc(ocm.OCM_LOCAL_HOST, n)
    g1, g2 = ocm.alloc(ocm.OCM_LOCAL_GPU, n), ocm.alloc(ocm.OCM_LOCAL_GPU, n)
    rg = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
    rh = ocm.alloc(ocm.OCM_REMOTE_HOST, n, n)
    ref = torch.randint(0, 256, (n,), dtype=torch.uint8)
    h1.local_tensor().copy_(ref)
    ocm.copy(g1, h1, n)                           # host -> GPU
    assert torch.equal(g1.local_tensor().cpu(), ref)
    ocm.copy(g2, g1, n)                           # GPU -> GPU (silent no-op in the reference, src/lib.c:635-658)
    assert torch.equal(g2.local_tensor().cpu(), ref)
    ocm.copy(rg, g2, n)                           # GPU -> remote GPU: window + far side in one kernel
    assert torch.equal(rg.local_tensor().cpu(), ref) and torch.equal(rg.remote_tensor().cpu(), ref)
    ocm.copy(h2, rg, n)                           # remote GPU -> host
    assert torch.equal(h2.local_tensor(), ref)
    h2.local_tensor().zero_()
    ocm.copy(rh, g1, n)                           # GPU -> remote host (pinned pool of the peer rank)
    assert torch.equal(rh.remote_tensor(), ref)
    ocm.copy(rg, rh, n)                           # remote host -> remote GPU (BUG(1) in the reference, src/lib.c:591-594)
    assert torch.equal(rg.remote_tensor().cpu(), ref)
    ocm.copy(rg, h1, n)                           # pinned host -> {GPU window, far GPU}: fused zero-copy read
    assert torch.equal(rg.remote_tensor().cpu(), ref)
    ocm.copy(h2, g2, n)                           # GPU -> host
    assert torch.equal(h2.local_tensor(), ref)
    # offsets on both hops
    ocm.copy(rg, g1, 4096, src_offset=8192, dest_offset=4096, src_offset_2=4096, dest_offset_2=1 << 20)
    assert torch.equal(rg.remote_tensor()[1 << 20:(1 << 20) + 4096].cpu(), ref[8192:8192 + 4096])
    for a in (h1, h2, g1, g2, rg, rh):
        a.free()


def test_copy_in_out_host_and_device_pointers(gpu_cluster):
    n = 1 << 20
    rg = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
    src_dev = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
    rg.copy_in(src_dev.data_ptr())
    assert torch.equal(rg.remote_tensor(), src_dev)
    out_host = torch.zeros(n, dtype=torch.uint8)
    rg.copy_out(out_host.data_ptr())
    assert torch.equal(out_host, src_dev.cpu())
    src_host = torch.randint(0, 256, (n,), dtype=torch.uint8)
    rg.copy_in(src_host.data_ptr())
    out_dev = torch.zeros(n, dtype=torch.uint8, device="cuda")
    rg.copy_out(out_dev.data_ptr())
    assert torch.equal(out_dev.cpu(), src_host)
    rg.free()


def test_async_stream_order_and_signal(gpu_cluster):
    n = 8 << 20
    a = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
    s = torch.cuda.Stream()
    w = a.local_tensor()
    with torch.cuda.stream(s):
        for i in range(4):
            ops.fill(w, i + 1, stream=s)
            a.put(n, stream=s)
    s.synchronize()
    a.wait_signal(4, timeout_ms=2000)
    assert a.signal() == 4
    assert ops.verify(a.remote_tensor(), 4, n) == 0
    a.free()


def test_freed_memory_is_scrubbed(gpu_cluster):
    import time

    n = 200 << 20     # window + far side: 400 MiB of the 1 GiB pool(s) per round -> blocks are reused quickly
    for it in range(8):
        a = None
        for _ in range(2000):                     # released blocks stay unavailable until the daemon zeroed them
            try:
                a = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
                break
            except ocm.OcmError:
                time.sleep(0.001)
        assert a is not None
        assert ops.verify(a.remote_tensor(), 0, n) == 0 and ops.verify(a.local_tensor(), 0, n) == 0
        ops.fill(a.local_tensor(), 0x1111111111111111 * (it + 1))
        torch.cuda.synchronize()
        a.put(n)
        a.free()


def test_strided_fp8_put_on_allocations(gpu_cluster):
    rows, elems = 256, 1024
    src = ocm.alloc(ocm.OCM_LOCAL_GPU, rows * 2048 * 4)
    x = src.local_tensor(torch.float32).view(rows, 2048)
    x.copy_(torch.randn(rows, 2048, device="cuda") * 3)
    payload, scales = rows * elems, rows * (elems // 32) * 4
    dst = ocm.alloc(ocm.OCM_REMOTE_GPU, 0, payload + scales, flags=ocm.ALLOC_NO_WINDOW)
    ocm.copy_strided(dst, src, row_bytes=elems * 4, counts=(rows, 1, 1), src_strides=(2048 * 4, 0, 0),
                     dst_strides=(elems, 0, 0), src_dtype=ocm.DT_F32, dst_dtype=ocm.DT_F8E4M3, scale_block=32,
                     scale_offset=payload)
    ocm.quiet()
    assert dst.signal() == 1
    back = ocm.alloc(ocm.OCM_LOCAL_GPU, rows * elems * 4)
    ocm.copy_strided(back, dst, row_bytes=elems, counts=(rows, 1, 1), src_strides=(elems, 0, 0),
                     dst_strides=(elems * 4, 0, 0), src_dtype=ocm.DT_F8E4M3, dst_dtype=ocm.DT_F32, scale_block=32,
                     scale_offset=payload)
    ocm.quiet()
    y = back.local_tensor(torch.float32).view(rows, elems)
    assert torch.allclose(y, x[:, :elems], rtol=0.0626, atol=1e-3)
    for a in (src, dst, back):
        a.free()


def test_striped_and_replicated_kinds(gpu_cluster):
    N = ocm.num_ranks()
    n = (6 << 20) + 4096
    ref = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
    # striped: one logical range block-distributed over every rank's pool
    s = ocm.alloc(ocm.OCM_STRIPED_GPU, n, n)
    assert s.kind == ocm.OCM_STRIPED_GPU and s.is_remote and s.remote_bytes >= n
    s.local_tensor().copy_(ref)
    torch.cuda.synchronize()
    s.put(n)
    per = s.remote_bytes // N
    for r in range(N):                               # every slab holds its block
        lo, hi = r * per, min(n, (r + 1) * per)
        if hi > lo:
            assert torch.equal(s.slab_tensor(r)[: hi - lo], ref[lo:hi])
    s.local_tensor().zero_()
    torch.cuda.synchronize()
    s.get(n)
    assert torch.equal(s.local_tensor(), ref)
    if N > 1:                                        # a window that straddles a stripe boundary
        s.local_tensor()[:8192].fill_(0xEE)
        torch.cuda.synchronize()
        s.put(8192, src_offset=0, dest_offset=per - 4096)
        assert int(s.slab_tensor(0)[per - 4096: per].sum()) == 0xEE * 4096
        assert int(s.slab_tensor(1)[:4096].sum()) == 0xEE * 4096
    # replicated: one put lands on every rank (cluster-multicast kernel), get reads the nearest replica
    rp = ocm.alloc(ocm.OCM_REPLICATED_GPU, n, n)
    assert rp.remote_bytes == n
    rp.local_tensor().copy_(ref)
    torch.cuda.synchronize()
    rp.put(n)
    for r in range(N):
        assert torch.equal(rp.slab_tensor(r)[:n], ref)
    rp.local_tensor().zero_()
    torch.cuda.synchronize()
    rp.get(n)
    assert torch.equal(rp.local_tensor(), ref)
    # two-sided copy from a plain GPU allocation into the striped range and back
    g = ocm.alloc(ocm.OCM_LOCAL_GPU, n)
    g.local_tensor().copy_(ref.flip(0))
    torch.cuda.synchronize()
    ocm.copy(s, g, n)
    g.local_tensor().zero_()
    torch.cuda.synchronize()
    ocm.copy(g, s, n)
    assert torch.equal(g.local_tensor(), ref.flip(0))
    for a in (s, rp, g):
        a.free()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="NVLS multicast needs two GPUs")
def test_nvls_multicast_replicated_put(gpu_cluster):
    """OCM_REPLICATED_GPU backed by the NVLS multicast region: one multimem.st stream lands on every GPU."""
    from oncilla_b200 import api

    if not api.multicast_ready(wait_s=10.0):
        pytest.skip("NVLS multicast not available on this box (daemons fell back to per-rank slabs)")
    N = ocm.num_ranks()
    n = (5 << 20) + 64
    ref = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
    a = ocm.alloc(ocm.OCM_REPLICATED_GPU, n, n)
    b = ocm.alloc(ocm.OCM_REPLICATED_GPU, 4096, 4096)      # a second range in the same region
    assert a.is_multicast and b.is_multicast
    for r in range(N):
        assert int(a.slab_tensor(r)[:n].sum()) == 0          # the region starts zeroed
    a.local_tensor().copy_(ref)
    torch.cuda.synchronize()
    before = ops.launch_count()
    a.put(n)
    assert ops.launch_count() == before + 1                  # one kernel for all replicas
    for r in range(N):
        assert torch.equal(a.slab_tensor(r)[:n], ref)
    assert int(b.slab_tensor(N - 1).sum()) == 0              # neighbours untouched
    # offsets and an unaligned tail fall back to per-replica stores but stay correct
    a.local_tensor()[:1000].fill_(7)
    torch.cuda.synchronize()
    a.put(1000, src_offset=0, dest_offset=8192)
    for r in range(N):
        assert int(a.slab_tensor(r)[8192:9192].sum()) == 7000
    a.local_tensor().zero_()
    torch.cuda.synchronize()
    a.get(n)
    assert torch.equal(a.local_tensor()[:8192], ref[:8192])
    a.free()
    c = ocm.alloc(ocm.OCM_REPLICATED_GPU, n, n)              # the range is reusable and was scrubbed
    assert c.is_multicast
    for r in range(N):
        assert int(c.slab_tensor(r)[:n].sum()) == 0
    c.free()
    b.free()


NVLS_WAITER = r"""
import sys, torch, oncilla_b200 as ocm
from oncilla_b200 import ops
ocm.init()
b = ocm.attach(bytes.fromhex(sys.argv[1]))
assert b.is_multicast and b.kind == ocm.OCM_REPLICATED_GPU
n = int(sys.argv[2])
print("WAITING", flush=True)
b.wait_signal(1, timeout_ms=20000)                      # rank 0's multicast put bumps OUR replica's word too
mine = b.slab_tensor(ocm.rank())
bad = ops.verify(mine[:n], 0xFEEDFACE12345678, n)
# and the other way: this application puts into the same range from its own GPU
src = ocm.alloc(ocm.OCM_LOCAL_GPU, n)
ops.fill(src.local_tensor(), 0x0BADF00D0BADF00D)
torch.cuda.synchronize()
ocm.copy(b, src, n)
sig = b.signal()
b.free(); src.free()
ocm.tini()
print("RESULT", bad, sig)
"""


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="NVLS multicast needs two GPUs")
def test_nvls_completion_signal_reaches_every_replica(gpu_cluster):
    """A multimem.st put ends with multimem.red on the allocation's completion word: an application on rank 1 that
    attached to the same range sees rank 0's put on ITS replica (ocm_signal_wait), and its own put counts here."""
    import subprocess
    import sys
    import time

    from conftest import REPO
    from oncilla_b200 import api

    if not api.multicast_ready(wait_s=10.0):
        pytest.skip("NVLS multicast not available on this box")
    n = 1 << 20
    a = ocm.alloc(ocm.OCM_REPLICATED_GPU, n, n)
    assert a.is_multicast and a.signal() == 0
    env = gpu_cluster.env(1)
    env["PYTHONPATH"] = str(REPO)
    p = subprocess.Popen([sys.executable, "-c", NVLS_WAITER, a.share().hex(), str(n)], env=env, stdout=subprocess.PIPE,
                         stderr=subprocess.PIPE, text=True)
    assert "WAITING" in p.stdout.readline()
    time.sleep(0.3)
    ops.fill(a.local_tensor(), 0xFEEDFACE12345678)
    torch.cuda.synchronize()
    a.put(n)
    assert a.signal() == 1
    out, err = p.communicate(timeout=120)
    assert p.returncode == 0 and "RESULT 0 2" in out, out[-1000:] + err[-2000:]
    a.wait_signal(2, timeout_ms=5000)
    assert ops.verify(a.slab_tensor(0)[:n], 0x0BADF00D0BADF00D, n) == 0      # rank 1's put landed in our replica
    # an unaligned put takes per-replica stores, and still counts once on every replica
    a.put(1000, src_offset=3, dest_offset=77)
    assert a.signal() == 3
    a.free()
    c = ocm.alloc(ocm.OCM_REPLICATED_GPU, n, n)                              # same range again: word back at zero
    assert c.is_multicast and c.signal() == 0
    c.free()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="NVLS multicast needs two GPUs")
def test_nvls_gather_reduce(gpu_cluster):
    """multimem.ld_reduce: the window receives the fp32 sum of every rank's replica, added inside the NVSwitch."""
    from oncilla_b200 import api

    if not api.multicast_ready(wait_s=10.0):
        pytest.skip("NVLS multicast not available on this box")
    N = ocm.num_ranks()
    elems = (3 << 18) + 4                                    # 3 MiB + 16 bytes of fp32
    nbytes = elems * 4
    a = ocm.alloc(ocm.OCM_REPLICATED_GPU, nbytes, nbytes)
    assert a.is_multicast
    g = torch.Generator(device="cuda").manual_seed(5)
    parts = [torch.randn(elems, device="cuda", generator=g) for _ in range(N)]
    for r in range(N):                                       # every replica gets a different contribution
        a.slab_tensor(r, torch.float32)[:elems].copy_(parts[r])
    torch.cuda.synchronize()
    a.local_tensor(torch.float32).fill_(-1.0)
    before = ops.launch_count()
    a.reduce_get_f32(nbytes, stream=torch.cuda.current_stream())
    torch.cuda.synchronize()
    assert ops.launch_count() == before + 1
    want = torch.stack(parts).sum(0)
    got = a.local_tensor(torch.float32)[:elems]
    assert torch.allclose(got, want, rtol=1e-6, atol=1e-6)
    # a sub-range at offsets
    a.local_tensor(torch.float32).fill_(-1.0)
    a.reduce_get_f32(4096, src_offset=1024, dest_offset=8192, stream=torch.cuda.current_stream())
    torch.cuda.synchronize()
    got = a.local_tensor(torch.float32)
    assert torch.allclose(got[256 : 256 + 1024], want[2048 : 2048 + 1024], rtol=1e-6, atol=1e-6)
    assert float(got[255]) == -1.0 and float(got[256 + 1024]) == -1.0
    with pytest.raises(ocm.OcmError):
        a.reduce_get_f32(4096, src_offset=8)                 # misaligned
    plain = ocm.alloc(ocm.OCM_REMOTE_GPU, 4096, 4096)
    with pytest.raises(ocm.OcmError):
        plain.reduce_get_f32(4096)                           # not a multicast allocation
    plain.free()
    a.free()


MC_LEAK = r"""
import os, torch, oncilla_b200 as ocm
from oncilla_b200 import api
ocm.init()
torch.cuda.set_device(ocm.device())
assert api.multicast_ready(wait_s=10.0)
a = ocm.alloc(ocm.OCM_REPLICATED_GPU, 1 << 20, 1 << 20)
assert a.is_multicast
a.local_tensor().fill_(0x5A); torch.cuda.synchronize(); a.put(1 << 20); torch.cuda.synchronize()
print("LEAKED", flush=True)
os._exit(0)                      # dies without ocm_free / ocm_tini
"""


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="NVLS multicast needs two GPUs")
def test_nvls_range_of_dead_application_is_reclaimed(gpu_cluster):
    """An application dies holding a multicast range full of data: the daemon of its rank marks the range dead,
    every daemon zeroes its own GPU's slice, the last one returns the range - the whole region is allocatable
    again and reads back zero on every GPU."""
    import time

    from oncilla_b200 import api

    if not api.multicast_ready(wait_s=10.0):
        pytest.skip("NVLS multicast not available on this box")
    N = ocm.num_ranks()
    r = run_py(gpu_cluster, N - 1, MC_LEAK)
    assert r.returncode == 0 and "LEAKED" in r.stdout, r.stdout + r.stderr
    # OCM_MC_BYTES default minus the granules that hold the region's completion words: only a fully free region fits
    region = (256 << 20) - (128 << 10)
    a, t0 = None, time.time()
    while time.time() - t0 < 15:
        a = ocm.alloc(ocm.OCM_REPLICATED_GPU, 4096, region)
        if a.is_multicast:
            break
        a.free()                                            # fell back to per-rank slabs: the range is not back yet
        a = None
        time.sleep(0.1)
    assert a is not None, "the dead application's multicast range was never reclaimed"
    for rk in range(N):
        assert int(a.slab_tensor(rk).sum()) == 0            # including the bytes the dead application wrote
    a.free()


ATTACH_PEER = r"""
import sys, torch
import oncilla_b200 as ocm
from oncilla_b200 import ops
ocm.init()
torch.cuda.set_device(ocm.device())
att = ocm.attach(bytes.fromhex(sys.argv[1]))
n = att.remote_bytes
src = ocm.alloc(ocm.OCM_LOCAL_GPU, n)
ops.fill(src.local_tensor(), 0xFEEDFACE12345678)
torch.cuda.synchronize()
ocm.copy(att, src, n // 2)                                   # two puts -> the owner's signal reaches 2
ocm.copy(att, src, n - n // 2, src_offset=n // 2, dest_offset=n // 2)
assert att.signal() == 2 and att.kind == ocm.OCM_REMOTE_GPU
att.free(); src.free(); ocm.tini()
print("PEER_OK")
"""


def test_attach_from_another_application(gpu_cluster):
    """A second application (on the last rank: another GPU when there is one) opens a block of this process
    by token and writes it with our put kernel; the owner waits on the block's completion signal."""
    n = (8 << 20) + 4096
    a = ocm.alloc(ocm.OCM_LOCAL_GPU, n)
    me = ocm.attach(a.share())                                # same process: aliases the same memory
    assert me.remote_ptr == a.local_ptr and a.attach_count == 1
    me.free()
    assert a.attach_count == 0
    r = run_py(gpu_cluster, gpu_cluster.n_ranks - 1, ATTACH_PEER, argv=[a.share().hex()])
    assert r.returncode == 0 and "PEER_OK" in r.stdout, r.stdout + r.stderr
    a.wait_signal(2, timeout_ms=10000)
    assert ops.verify(a.local_tensor(), 0xFEEDFACE12345678, n) == 0
    assert a.attach_count == 0
    a.free()


def test_native_ocm_test_program_gpu(gpu_cluster):
    for args in (["1", "4", "4", "2"], ["1", "4", "8", "7"], ["2", "4", "4", "7"], ["3", "2", "2", "7"],
                 ["4", "1", "1", "7", "3"], ["5", "4", "4", "2"], ["5", "4", "4", "7"]):
        r = run_app(gpu_cluster, 0, [BIN / "ocm_test", *args])
        assert r.returncode == 0, r.stdout + r.stderr
        assert "pass: OCM test completed successfully" in r.stdout


def test_batched_onesided_copies_one_launch(gpu_cluster):
    """ocm_copy_onesided_batch: 100 scattered small puts (odd sizes and alignments among them) cost 2 launches
    (64 segments per launch), land byte-exact, and the signal counts every put; gets come back the same way."""
    n_msg, span = 100, 1 << 20
    a = ocm.alloc(ocm.OCM_REMOTE_GPU, span, span)
    g = torch.Generator().manual_seed(11)
    ref = torch.randint(0, 256, (span,), dtype=torch.uint8, generator=g)
    a.local_tensor().copy_(ref.cuda())
    torch.cuda.synchronize()
    entries, expect = [], torch.zeros(span, dtype=torch.uint8)
    slot = span // n_msg
    for i in range(n_msg):
        nbytes = int(torch.randint(1, slot - 64, (1,), generator=g))
        so = i * slot + int(torch.randint(0, 32, (1,), generator=g))
        do = i * slot + (so % 32 if i % 3 else int(torch.randint(0, 32, (1,), generator=g)))   # some congruent mod 32, some not
        entries.append(("put", nbytes, so, do))
        expect[do:do + nbytes] = ref[so:so + nbytes]
    lc = ops.launch_count()
    a.batch(entries)
    ocm.quiet()
    assert ops.launch_count() - lc == 2
    assert a.signal() == n_msg
    assert torch.equal(a.remote_tensor().cpu(), expect)
    # and back: window <- far side, swapped offsets
    a.local_tensor().zero_()
    torch.cuda.synchronize()
    a.batch([("get", nb, do, so) for (_, nb, so, do) in entries])        # window[do..] <- far[so..]
    ocm.quiet()
    back = a.local_tensor().cpu()
    for (_, nb, so, do) in entries:
        assert torch.equal(back[do:do + nb], expect[so:so + nb])
    assert a.signal() == n_msg                                           # gets do not signal
    with pytest.raises(ocm.OcmError):
        a.batch([("put", 16, span - 8, 0)])                              # out of range: refused as a whole
    a.free()


def test_resident_copy_server_serves_small_blocking_copies(gpu_cluster):
    """H2: blocking copies up to OCM_SERVER_MAX bytes are moved by the resident kernel (no launch); the kernel leaves
    when idle and comes back on demand; work queued asynchronously on the library stream is never overtaken."""
    import time

    s0 = ocm.server_stats()
    assert s0["enabled"]
    n = 48 << 10
    a = ocm.alloc(ocm.OCM_REMOTE_GPU, 4 << 20, 4 << 20)
    w = a.local_tensor()
    launches0 = ops.launch_count()
    for i in range(20):
        ops.fill(w[i * n:(i + 1) * n], 0xABCD000000000000 | i)
    torch.cuda.synchronize()
    lc = ops.launch_count()
    for i in range(20):                                                  # back to back: the kernel stays resident
        a.put(n, src_offset=i * n, dest_offset=(19 - i) * n)             # blocking: through the server
        assert a.signal() == i + 1
    assert ops.launch_count() - lc <= 2, "kernels were launched for small blocking puts"
    for i in range(20):
        assert ops.verify(a.remote_tensor()[(19 - i) * n:(20 - i) * n], 0xABCD000000000000 | i, n) == 0
    s1 = ocm.server_stats()
    assert s1["ops"] - s0["ops"] == 20 and 1 <= s1["launches"] - s0["launches"] <= 2
    # unaligned pieces and odd sizes take the bytewise path of the server
    h = ocm.alloc(ocm.OCM_LOCAL_HOST, 1 << 16)
    ref = torch.arange(1 << 16, dtype=torch.int32).to(torch.uint8)
    h.local_tensor().copy_(ref)
    ocm.copy(a, h, 1001, src_offset=3, dest_offset=7, src_offset_2=7, dest_offset_2=13)
    assert torch.equal(a.local_tensor()[7:7 + 1001].cpu(), ref[3:3 + 1001])
    assert torch.equal(a.remote_tensor()[13:13 + 1001].cpu(), ref[3:3 + 1001])
    # the kernel is resident on demand only: after an idle spell it has left (synchronize returns) and is restarted
    time.sleep(0.05)
    t0 = time.perf_counter()
    torch.cuda.synchronize()
    assert time.perf_counter() - t0 < 0.05
    a.get(n)
    s2 = ocm.server_stats()
    assert s2["launches"] > s1["launches"] and s2["ops"] > s1["ops"]
    # stream order: a big asynchronous put on the library stream, then a small blocking get of its last bytes
    big = 4 << 20
    ops.fill(w, 0x1122334455667788)
    torch.cuda.synchronize()
    a.put(big, stream=None)                                                # async, library stream
    ops.fill(w[:4096], 0)
    torch.cuda.synchronize()
    a.get(4096, src_offset=0, dest_offset=big - 4096)                      # must queue behind the put
    assert ops.verify(w[:4096], 0x1122334455667788, 4096) == 0
    ocm.quiet()
    assert ops.launch_count() > launches0
    a.free()
    h.free()


def test_blocking_copies_from_many_threads_gpu(gpu_cluster):
    """eight threads of one process issue blocking GPU copies at once; each waits on its OWN
    completion slot, so no call may return before its own kernel has run (tests/cpp/ocm_mt.c checks every byte
    right after each call)."""
    r = run_app(gpu_cluster, 0, [BIN / "ocm_mt", "8", "150", "gpu"], timeout=300)
    assert r.returncode == 0 and "pass: ocm_mt 8 threads x 150 iterations" in r.stdout and "FAIL" not in r.stdout, \
        r.stdout[-2000:] + r.stderr[-2000:]


def test_examples_on_gpu(gpu_cluster):
    """examples/: the C hello program with a remote-GPU pair, and a user kernel that works on the far side
    through ocm_remote_ptr()."""
    r = run_app(gpu_cluster, 0, [BIN / "hello_ocm", "7"])
    assert r.returncode == 0 and "pass: hello_ocm" in r.stdout, r.stdout + r.stderr
    r = run_app(gpu_cluster, 0, [BIN / "peer_kernel"])
    assert r.returncode == 0 and "pass: peer_kernel" in r.stdout, r.stdout + r.stderr


def test_workloads_gups_and_sort(gpu_cluster):
    from oncilla_b200.models import gups, sort

    res = gups.run(total_words=1 << 22, n_updates=1 << 22, iters=2)
    assert res["verified"] is True and res["gups"] > 0
    res = gups.run(total_words=1 << 22, n_updates=1 << 22, iters=1, local_only=True)
    assert res["verified"] is True and res["n_ranks"] == 1
    res = sort.run(total_keys=1 << 20, chunk_keys=1 << 18)
    assert res["ok"] and res["n_runs"] == 4 and res["final"] == "merge"
    res = sort.run(total_keys=(1 << 20) + 777, chunk_keys=1 << 18, final="sort")
    assert res["ok"] and res["n_runs"] == 5
    # force the host tier: HBM budget of one run per rank
    res = sort.run(total_keys=1 << 20, chunk_keys=1 << 18, gpu_budget_per_rank=(1 << 18) * 8)
    assert res["ok"] and res["bytes_in_host"] > 0


def test_distributed_sort_single_rank_and_torchrun(gpu_cluster):
    """The sample sort over shared blocks: in this process (world 1), then as a torchrun job with one
    application per GPU (2 ranks when there are two GPUs) that exchanges keys with one-sided puts."""
    import os
    import sys

    from conftest import REPO
    from oncilla_b200.models import dsort

    res = dsort.run(keys_per_rank=1 << 18)
    assert res["ok"] and res["recv_keys"] == 1 << 18
    n = min(torch.cuda.device_count(), 2)
    env = {k: v for k, v in os.environ.items() if not k.startswith("OCM_")}
    env["PYTHONPATH"] = str(REPO) + os.pathsep + env.get("PYTHONPATH", "")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", "29631", str(REPO / "bench" / "workloads.py"),
                        "--only", "dsort", "--dsort-keys", str(1 << 20)], env=env, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if '"distributed_sort"' in l][-1]
    import json

    out = json.loads(line)
    assert out["verified"] is True and out["n_gpus"] == n
    # the out-of-core pipeline as a job of n ranks: runs in everybody's pools, slices fetched through attached handles,
    # half of the data forced onto the pinned host tier
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", "29633", str(REPO / "bench" / "workloads.py"),
                        "--only", "oocsort", "--sort-keys", str(1 << 21), "--sort-chunk", str(1 << 19), "--sort-final", "both",
                        "--gpu-bytes", "1G", "--host-bytes", "512M"], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if '"ooc_sort"' in l]
    assert [l["final"] for l in lines] == ["merge", "sort"] and all(l["verified"] is True for l in lines), lines
    assert lines[0]["bytes_in_host_rank0"] > 0 and lines[0]["slices_per_partition"] == 4 * n


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_gpu_is_really_remote(gpu_cluster):
    n = 64 << 20
    a = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
    assert a.remote_rank == 1
    ops.fill(a.local_tensor(), 0xABCDEF)
    torch.cuda.synchronize()
    a.put(n)
    # ask a process attached to rank 1 (GPU 1) to checksum its own pool region through its daemon's view
    code = f"""
import torch, oncilla_b200 as ocm
from oncilla_b200 import ops
ocm.init()
assert ocm.device() == 1                        # attached to rank 1 -> GPU 1; torch's current device stays 0 on purpose
b = ocm.alloc(ocm.OCM_REMOTE_GPU, {n}, {n})     # lands on rank 0 (neighbour of 1): other direction
ops.fill(b.local_tensor(), 0x55)
torch.cuda.synchronize(1)
b.put({n}); b.get({n})
print("BAD", ops.verify(b.local_tensor(), 0x55))
b.free(); ocm.tini()
"""
    r = run_py(gpu_cluster, 1, code)
    assert r.returncode == 0 and "BAD 0" in r.stdout, r.stdout + r.stderr
    assert ops.verify(a.remote_tensor(), 0xABCDEF, n) == 0
    a.free()
