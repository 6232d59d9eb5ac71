"""Numerics of every sm_100a kernel against a plain fp32 / byte-exact reference of the same op.

References: the C++ oracles in src/oracle (through the C ABI) and plain PyTorch expressions.
The reference project has no kernels to compare with; its copy semantics are memcpy with
the same offsets (reference src/lib.c:523), which is what the copy tests assert.
"""
import ctypes

import pytest
import torch

import oncilla_b200 as ocm
from oncilla_b200 import ops

pytestmark = pytest.mark.gpu


def rnd_bytes(n, seed=0):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    return torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda", generator=g)


@pytest.mark.parametrize("mode", [None, 1, 2])  # auto (256-bit vector), 1 bulk (TMA), 2 vector (16 B)
@pytest.mark.parametrize("nbytes", [16, 4096, 65536 + 16, (1 << 20) + 48, (64 << 20) + 7])
def test_copy_matches_memcpy(mode, nbytes):
    src = rnd_bytes(nbytes + 64, seed=nbytes)
    dst = torch.zeros(nbytes + 64, dtype=torch.uint8, device="cuda")
    ops.copy(dst, src, nbytes, mode=mode)
    torch.cuda.synchronize()
    assert torch.equal(dst[:nbytes], src[:nbytes])
    assert int(dst[nbytes:].sum()) == 0          # nothing written past the end


@pytest.mark.parametrize("src_off,dst_off", [(0, 0), (16, 48), (8, 8), (4, 12), (1, 3), (3, 3)])
def test_copy_alignments(src_off, dst_off):
    n = (1 << 20) + 13
    src = rnd_bytes(n + 64, seed=7)
    dst = torch.zeros(n + 128, dtype=torch.uint8, device="cuda")
    ops.copy(dst[dst_off:], src[src_off:], n)
    torch.cuda.synchronize()
    assert torch.equal(dst[dst_off:dst_off + n], src[src_off:src_off + n])
    assert int(dst[:dst_off].sum()) == 0 and int(dst[dst_off + n:].sum()) == 0


@pytest.mark.parametrize("mode", [None, 1, 2])
def test_copy_two_destinations_and_signal(mode):
    n = (8 << 20) + 32
    src = rnd_bytes(n, seed=3)
    d1 = torch.zeros(n, dtype=torch.uint8, device="cuda")
    d2 = torch.zeros(n, dtype=torch.uint8, device="cuda")
    sig = torch.zeros(1, dtype=torch.int64, device="cuda")
    for k in range(3):
        ops.copy(d1, src, n, dst2=d2, signal=sig, mode=mode)
    torch.cuda.synchronize()
    assert torch.equal(d1, src) and torch.equal(d2, src)
    assert int(sig.item()) == 3                  # exactly one increment per put, many CTAs


@pytest.mark.parametrize("ctas,agents,stages,tile", [(1, 1, 2, 4096), (8, 4, 3, 16384), (148, 8, 2, 8192),
                                                     (32, 2, 6, 16384), (16, 1, 3, 65536)])
def test_copy_tunings(ctas, agents, stages, tile):
    n = (16 << 20) + 16 * 5
    src = rnd_bytes(n, seed=11)
    dst = torch.zeros(n, dtype=torch.uint8, device="cuda")
    ops.copy(dst, src, n, mode=1, ctas=ctas, agents=agents, stages=stages, tile_bytes=tile)
    torch.cuda.synchronize()
    assert torch.equal(dst, src)


def test_copy_pinned_host_zero_copy():
    n = (4 << 20) + 16
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    host.copy_(torch.arange(n, dtype=torch.int64).to(torch.uint8))
    dst = torch.zeros(n, dtype=torch.uint8, device="cuda")
    ops.copy(dst, host, n, mode=1)                # TMA reads pinned host memory directly
    torch.cuda.synchronize()
    assert torch.equal(dst.cpu(), host)
    back = torch.zeros(n, dtype=torch.uint8).pin_memory()
    ops.copy(back, dst, n, mode=2)                # kernel stores into pinned host memory
    torch.cuda.synchronize()
    assert torch.equal(back, host)


@pytest.mark.parametrize("nbytes,off", [(1, 0), (15, 1), (4096, 0), ((1 << 20) + 5, 3), (32 << 20, 0)])
def test_fill_verify_checksum(nbytes, off):
    import ctypes

    buf = torch.full((nbytes + 64,), 0xAA, dtype=torch.uint8, device="cuda")
    pat = 0x0102030405060708
    ops.fill(buf[off:], pat, nbytes)
    torch.cuda.synchronize()
    host = buf.cpu()
    ref = torch.full((nbytes + 64,), 0xAA, dtype=torch.uint8)
    ocm._native.lib.ocm_oracle_fill(ctypes.c_void_p(ref[off:].data_ptr()), nbytes, ctypes.c_uint64(pat))
    assert torch.equal(host, ref)
    assert ops.verify(buf[off:], pat, nbytes) == 0
    buf[off + nbytes // 2] ^= 0xFF
    assert ops.verify(buf[off:], pat, nbytes) == 1
    host2 = buf.cpu()                            # keep the host copy alive across the C call
    want = ocm._native.lib.ocm_oracle_checksum(ctypes.c_void_p(host2[off:].data_ptr()), nbytes)
    assert ops.checksum(buf[off:], nbytes) == want


@pytest.mark.parametrize("rows,row_bytes,src_pitch,dst_pitch", [
    (512, 4096, 8192, 4096),        # pack: bulk rows, several rows per stage
    (64, 65536, 70000 // 16 * 16, 65536),   # long rows split into chunks
    (1000, 48, 64, 48),             # short rows: vector path
    (300, 100, 107, 101),           # nothing aligned: byte path
    (256, 1024, 1024, 2048),        # unpack (scatter)
])
def test_strided_bytes(rows, row_bytes, src_pitch, dst_pitch):
    src = rnd_bytes(rows * src_pitch + 64, seed=rows)
    dst = torch.zeros(rows * dst_pitch + 64, dtype=torch.uint8, device="cuda")
    ops.strided_copy(dst, src, row_bytes=row_bytes, counts=(rows, 1, 1), src_strides=(src_pitch, 0, 0),
                     dst_strides=(dst_pitch, 0, 0))
    torch.cuda.synchronize()
    s = src[:rows * src_pitch].view(rows, src_pitch)[:, :row_bytes]
    d = dst[:rows * dst_pitch].view(rows, dst_pitch)
    assert torch.equal(d[:, :row_bytes], s)
    if dst_pitch > row_bytes:
        assert int(d[:, row_bytes:].sum()) == 0


def test_strided_3d_box():
    # a (6, 10, 2048-byte) box out of a (8, 16, 4096-byte) array into a dense destination
    src = rnd_bytes(8 * 16 * 4096, seed=5).view(8, 16, 4096)
    dst = torch.zeros(6 * 10 * 2048, dtype=torch.uint8, device="cuda")
    ops.strided_copy(dst, src[1:, 3:, 1024:], row_bytes=2048, counts=(10, 6, 1),
                     src_strides=(4096, 16 * 4096, 0), dst_strides=(2048, 10 * 2048, 0))
    torch.cuda.synchronize()
    assert torch.equal(dst.view(6, 10, 2048), src[1:7, 3:13, 1024:3072])


STRIDED_ENGINES_SNIPPET = r"""
import sys, torch, oncilla_b200 as ocm
from oncilla_b200 import ops
torch.manual_seed(3)
def check(rows, row_bytes, sp, dp, off_s=0, off_d=0):
    src = torch.randint(0, 256, (rows * sp + 256,), dtype=torch.uint8, device="cuda")
    dst = torch.zeros(rows * dp + 256, dtype=torch.uint8, device="cuda")
    ops.strided_copy(dst[off_d:], src[off_s:], row_bytes=row_bytes, counts=(rows, 1, 1), src_strides=(sp, 0, 0), dst_strides=(dp, 0, 0))
    torch.cuda.synchronize()
    s = src[off_s:off_s + rows * sp].view(rows, sp)[:, :row_bytes]
    d = dst[off_d:off_d + rows * dp].view(rows, dp)
    assert torch.equal(d[:, :row_bytes], s), (rows, row_bytes, sp, dp)
    if dp > row_bytes:
        assert int(d[:, row_bytes:].sum()) == 0, "bytes outside the rows were written"
    assert int(dst[:off_d].sum()) == 0 and int(dst[off_d + rows * dp:].sum()) == 0
for args in [(512, 4096, 8192, 4096), (4096, 4096, 8192, 4096), (64, 65536, 69632, 65536), (777, 2048 + 64, 4096, 2048 + 64),
             (5000, 256, 512, 256), (3000, 64, 128, 64), (1000, 48, 64, 48), (256, 1024, 1024, 2048), (33, 16, 32, 16),
             (1500, 4096 + 16, 8192 + 16, 4096 + 16, 16, 32), (2, 1 << 20, 1 << 21, 1 << 20), (300, 100, 107, 101)]:
    check(*args)
# 3-D box out of a (8, 16, 4096-byte) array into a dense destination, and scattered back into a zeroed array
src = torch.randint(0, 256, (8, 16, 4096), dtype=torch.uint8, device="cuda")
dst = torch.zeros(6 * 10 * 2048, dtype=torch.uint8, device="cuda")
ops.strided_copy(dst, src[1:, 3:, 1024:], row_bytes=2048, counts=(10, 6, 1), src_strides=(4096, 16 * 4096, 0), dst_strides=(2048, 10 * 2048, 0))
back = torch.zeros_like(src)
ops.strided_copy(back[1:, 3:, 1024:], dst, row_bytes=2048, counts=(10, 6, 1), src_strides=(2048, 10 * 2048, 0), dst_strides=(4096, 16 * 4096, 0))
torch.cuda.synchronize()
assert torch.equal(dst.view(6, 10, 2048), src[1:7, 3:13, 1024:3072])
want = torch.zeros_like(src); want[1:7, 3:13, 1024:3072] = src[1:7, 3:13, 1024:3072]
assert torch.equal(back, want)
# 4-D: (3, 4, 5) rows of 512 bytes
src4 = torch.randint(0, 256, (3, 4, 5, 1024), dtype=torch.uint8, device="cuda")
d4 = torch.zeros(3, 4, 5, 512, dtype=torch.uint8, device="cuda")
ops.strided_copy(d4, src4[..., 256:], row_bytes=512, counts=(5, 4, 3), src_strides=(1024, 5 * 1024, 20 * 1024), dst_strides=(512, 5 * 512, 20 * 512))
torch.cuda.synchronize()
assert torch.equal(d4, src4[..., 256:768])
print("ENGINE OK", ops.launch_count())
"""


@pytest.mark.parametrize("engine", ["auto", "tma", "wide", "bulk", "vec"])
def test_strided_engines_agree(engine):
    """Every strided engine (TMA tensor maps, wide vectors, 1-D bulk rows, narrow vectors) moves the same boxes byte for
    byte, edge tiles and unaligned fall-backs included.  The engine is chosen once per process (OCM_STRIDED_MODE)."""
    import os
    import subprocess
    import sys

    from conftest import REPO

    env = dict(os.environ, OCM_STRIDED_MODE=engine, PYTHONPATH=str(REPO))
    r = subprocess.run([sys.executable, "-c", STRIDED_ENGINES_SNIPPET], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ENGINE OK" in r.stdout, r.stdout[-1500:] + r.stderr[-3000:]


def _oracle_strided(dst_cpu, src_cpu, **kw):
    ops.oracle_strided_copy(dst_cpu, src_cpu, **kw)


@pytest.mark.parametrize("src_dt,elems", [(ocm.DT_F32, 256), (ocm.DT_BF16, 384), (ocm.DT_F32, 200), (ocm.DT_BF16, 100)])
@pytest.mark.parametrize("blk", [32, 128])
def test_cast_fp8_block_scaled_matches_oracle(src_dt, elems, blk):
    rows, pitch_e = 96, 512
    tdt = torch.float32 if src_dt == ocm.DT_F32 else torch.bfloat16
    es = 4 if src_dt == ocm.DT_F32 else 2
    g = torch.Generator(device="cuda")
    g.manual_seed(elems * blk)
    x = (torch.randn(rows, pitch_e, device="cuda", generator=g) * torch.logspace(-3, 3, rows, device="cuda")[:, None]).to(tdt)
    x[5, :64] = 0                                   # an all-zero block -> scale 1
    bpr = (elems + blk - 1) // blk
    q = torch.zeros(rows, elems, dtype=torch.uint8, device="cuda")
    sc = torch.zeros(rows, bpr, dtype=torch.float32, device="cuda")
    kw = dict(row_bytes=elems * es, counts=(rows, 1, 1), src_strides=(pitch_e * es, 0, 0), dst_strides=(elems, 0, 0),
              src_dtype=src_dt, dst_dtype=ocm.DT_F8E4M3, scale_block=blk)
    ops.strided_copy(q, x, scales=sc, **kw)
    torch.cuda.synchronize()
    xc, qc, scc = x.cpu(), torch.zeros(rows, elems, dtype=torch.uint8), torch.zeros(rows, bpr, dtype=torch.float32)
    _oracle_strided(qc, xc, scales=scc, **kw)
    assert torch.equal(sc.cpu(), scc)               # scales bit-exact (IEEE division on both sides)
    assert torch.equal(q.cpu(), qc)                 # payload bit-exact (RNE, satfinite)
    # and back: dequantise on the GPU, compare with the oracle and with the fp32 source
    y = torch.zeros(rows, elems, dtype=tdt, device="cuda")
    kw2 = dict(row_bytes=elems, counts=(rows, 1, 1), src_strides=(elems, 0, 0), dst_strides=(elems * es, 0, 0),
               src_dtype=ocm.DT_F8E4M3, dst_dtype=src_dt, scale_block=blk)
    ops.strided_copy(y, q, scales=sc, **kw2)
    torch.cuda.synchronize()
    yc = torch.zeros(rows, elems, dtype=tdt)
    _oracle_strided(yc, qc, scales=scc, **kw2)
    assert torch.equal(y.cpu().view(torch.uint8), yc.view(torch.uint8))
    ref = x[:, :elems].float()
    err = (y.float() - ref).abs()
    tol = ref.abs() * 0.0625 + sc.repeat_interleave(blk, dim=1)[:, :elems] * 2.0 ** -9 + 1e-30
    if src_dt == ocm.DT_BF16:
        tol = tol * 1.02 + ref.abs() * 2.0 ** -8
    assert bool((err <= tol).all())


def test_cast_f32_bf16_roundtrip_matches_torch():
    x = torch.randn(128, 1024, device="cuda")
    y = torch.zeros(128, 1024, dtype=torch.bfloat16, device="cuda")
    ops.strided_copy(y, x, row_bytes=4096, counts=(128, 1, 1), src_strides=(4096, 0, 0), dst_strides=(2048, 0, 0),
                     src_dtype=ocm.DT_F32, dst_dtype=ocm.DT_BF16)
    z = torch.zeros_like(x)
    ops.strided_copy(z, y, row_bytes=2048, counts=(128, 1, 1), src_strides=(2048, 0, 0), dst_strides=(4096, 0, 0),
                     src_dtype=ocm.DT_BF16, dst_dtype=ocm.DT_F32)
    torch.cuda.synchronize()
    assert torch.equal(y, x.to(torch.bfloat16))     # same rounding as torch (RNE)
    assert torch.equal(z, x.to(torch.bfloat16).float())


@pytest.mark.parametrize("n_dst", [2, 4, 8])
def test_put_replicated_cluster_kernel(n_dst):
    """K6 through the raw-pointer ABI: one CTA of a cluster per destination, every CTA bulk-loads its own tiles (the
    cluster-shared form needs a source the library knows to be local: test_gpu_resilience.py, replicated put
    without NVLS)."""
    n = (6 << 20) + 16
    src = rnd_bytes(n, seed=n_dst)
    dsts = [torch.zeros(n, dtype=torch.uint8, device="cuda") for _ in range(n_dst)]
    ops.put_replicated(dsts, src, n)
    torch.cuda.synchronize()
    for d in dsts:
        assert torch.equal(d, src)


def test_gups_matches_oracle_and_is_involution():
    import ctypes

    total, wps, n_upd, seed = 1 << 16, 1 << 10, 1 << 18, 42
    slabs = [torch.zeros(total // 4, dtype=torch.int64, device="cuda") for _ in range(4)]
    ops.gups(slabs, wps, total, n_upd, seed)
    torch.cuda.synchronize()
    ref = torch.zeros(total, dtype=torch.int64)
    ocm._native.lib.ocm_oracle_gups(ctypes.c_void_p(ref.data_ptr()), total, n_upd, seed, ops.gups_threads(n_upd))
    # un-stripe the device table: word i lives on slab (i // wps) % 4
    got = torch.stack([s.cpu().view(-1, wps) for s in slabs], dim=1).reshape(-1)
    assert torch.equal(got, ref)
    ops.gups(slabs, wps, total, n_upd, seed)
    torch.cuda.synchronize()
    assert all(int(s.abs().sum()) == 0 for s in slabs)


@pytest.mark.parametrize("n", [1, 2, 1000, 2048, 2049, 8192, 100_003, 1 << 19, (1 << 19) + 1, (1 << 20) + 17])
def test_sort_u64_matches_oracle(n):
    """Sizes on both sides of the fused small-input mode (<= 2^19 keys: the scatter kernel of a pass builds the
    next pass's histogram) and of the one-tile-per-CTA boundary."""
    g = torch.Generator(device="cuda")
    g.manual_seed(n)
    # the full 64-bit range: about half of the keys have bit 63 set (negative as int64, LARGEST as unsigned keys)
    keys = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
    keys[: n // 3] = keys[: n // 3] & 0xFFFF        # plenty of duplicates and small keys
    keys[n // 2:n // 2 + 3] = torch.tensor([-1, -2**63, 2**63 - 1], dtype=torch.int64, device="cuda")[: max(0, min(3, n - n // 2))]
    # unsigned order = signed order after flipping the sign bit
    flip = torch.tensor(-2**63, dtype=torch.int64, device="cuda")
    want = torch.sort(keys ^ flip).values ^ flip
    host = keys.cpu().clone()
    ops.sort_u64(keys)
    torch.cuda.synchronize()
    assert torch.equal(keys, want)
    if n <= 100_003:                                 # and against the plain C++ oracle (std::sort on uint64_t)
        ocm._native.lib.ocm_oracle_sort_u64(ctypes.c_void_p(host.data_ptr()), host.numel())
        assert torch.equal(keys.cpu(), host)


@pytest.mark.parametrize("k,total", [(1, 1000), (2, 5), (2, 100_000), (3, 4096), (7, 300_001), (16, 1 << 20), (33, 2_000_003)])
def test_merge_runs_matches_sorted_concatenation(k, total):
    """k sorted runs of uneven length (an empty one among them) merge into the sorted whole; full-range unsigned keys
    with duplicates, tile edges (2048 outputs per tile) on both sides."""
    g = torch.Generator(device="cuda")
    g.manual_seed(k * 1000 + total % 997)
    flip = torch.tensor(-2**63, dtype=torch.int64, device="cuda")
    cuts = sorted(int(x) for x in torch.randint(0, total + 1, (k - 1,), generator=torch.Generator().manual_seed(k)))
    if k > 2:
        cuts[1] = cuts[0]                                   # an empty run
    offs = [0] + cuts + [total]
    keys = torch.randint(-2**63, 2**63 - 1, (total,), dtype=torch.int64, device="cuda", generator=g)
    keys[: total // 5] &= 0xFF                             # many equal keys across runs
    keys = keys[torch.randperm(total, device="cuda", generator=g)]
    for lo, hi in zip(offs[:-1], offs[1:]):
        keys[lo:hi] = torch.sort(keys[lo:hi] ^ flip).values ^ flip
    want = torch.sort(keys ^ flip).values ^ flip
    out = ops.merge_runs(keys, torch.empty_like(keys), offs)
    torch.cuda.synchronize()
    assert torch.equal(out, want)


def test_launch_counter_counts_our_kernels():
    before = ops.launch_count()
    a, b = rnd_bytes(1 << 20), torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    ops.copy(b, a)
    ops.fill(b, 0)
    torch.cuda.synchronize()
    assert ops.launch_count() == before + 2
