"""The allocation-free sort entry point, checked against the regular one.

`ocmk_sort_u64_noalloc` exists for callers that cannot obtain device memory at call time: an
application running next to daemons that already own almost all of the HBM, or code in a
latency-critical section where cudaMalloc/cudaFree could stall the whole device.  It keeps its
scratch in a fixed array inside the library (at most OCMK_SORT_NOALLOC_MAX_KEYS keys per call).

This file is named so that pytest collects it last: if it fails under `-x`, every other test has
already run.  The test goes through the C ABI only, makes two calls in a row (the scratch keeps the
first call's contents), and must pass alone in a fresh process as well as after the other tests.
"""
import ctypes

import pytest
import torch

import oncilla_b200 as ocm

pytestmark = pytest.mark.gpu

MAX_KEYS = 1 << 24
lib = ocm._native.lib


def _keys(n, seed):
    g = torch.Generator(device="cuda")
    g.manual_seed(seed)
    # full 64-bit range: keys with bit 63 set are negative as int64 but the LARGEST as unsigned sort keys
    k = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
    k[: n // 4] &= 0xFFFFF          # duplicates and short keys exercise every digit position
    return k


def _unsigned_sorted(keys):
    flip = torch.tensor(-2**63, dtype=torch.int64, device=keys.device)
    return torch.sort(keys ^ flip).values ^ flip


def _sort(fn, keys):
    work, tmp = keys.clone(), torch.empty_like(keys)
    stream = ctypes.c_void_p(int(torch.cuda.current_stream().cuda_stream))
    rc = fn(ctypes.c_void_p(work.data_ptr()), ctypes.c_void_p(tmp.data_ptr()), keys.numel(), stream)
    torch.cuda.synchronize()
    return rc, work


def test_noalloc_sort_matches_regular_entry_point_twice():
    torch.cuda.set_device(0)
    for call, (n, seed) in enumerate([((1 << 20) + 12345, 1), (300_001, 2)]):   # two calls in a row, different sizes
        keys = _keys(n, seed)
        rc_ref, want = _sort(lib.ocmk_sort_u64, keys)
        rc_new, got = _sort(lib.ocmk_sort_u64_noalloc, keys)
        assert rc_ref == 0 and rc_new == 0, (call, rc_ref, rc_new)
        assert torch.equal(got, want), f"call {call}: outputs differ"        # exact: same tolerance as the sort tests
        assert torch.equal(got, _unsigned_sorted(keys))
        host = keys.cpu().clone()                                             # and the plain C++ oracle
        lib.ocm_oracle_sort_u64(ctypes.c_void_p(host.data_ptr()), host.numel())
        assert torch.equal(got.cpu(), host)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs a second GPU")
def test_noalloc_sort_on_a_second_device():
    """The scratch array exists once per GPU and its address is looked up per device: a call on GPU 1 (first use
    there) must not reuse GPU 0's address."""
    torch.cuda.set_device(0)
    _sort(lib.ocmk_sort_u64_noalloc, _keys(50_000, 5))                        # make sure GPU 0's address is cached first
    try:
        torch.cuda.set_device(1)
        keys = _keys(200_000, 9).to("cuda:1")
        work, tmp = keys.clone(), torch.empty_like(keys)
        rc = lib.ocmk_sort_u64_noalloc(ctypes.c_void_p(work.data_ptr()), ctypes.c_void_p(tmp.data_ptr()), keys.numel(),
                                       ctypes.c_void_p(int(torch.cuda.current_stream().cuda_stream)))
        torch.cuda.synchronize()
        assert rc == 0 and torch.equal(work, _unsigned_sorted(keys))
    finally:
        torch.cuda.set_device(0)


def test_noalloc_sort_rejects_oversize_without_launching():
    torch.cuda.set_device(0)
    keys = torch.zeros(8, dtype=torch.int64, device="cuda")
    before = keys.clone()
    rc = lib.ocmk_sort_u64_noalloc(ctypes.c_void_p(keys.data_ptr()), ctypes.c_void_p(keys.data_ptr()), MAX_KEYS + 1,
                                   ctypes.c_void_p(0))
    torch.cuda.synchronize()
    assert rc == -5                                   # -OCM_ERANGE
    assert torch.equal(keys, before)
