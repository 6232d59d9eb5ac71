"""Torch-free check of the Python binding on one GPU (starts in a second: nothing imports torch): sync and
stream-ordered one-sided calls on a remote-GPU pair and on a host pair, signals, whole-allocation copy_out."""
import ctypes as C
import sys

sys.path.insert(0, __file__.rsplit("/scripts/", 1)[0])
import oncilla_b200 as ocm  # noqa: E402
from oncilla_b200 import _native as N  # noqa: E402

assert "torch" not in sys.modules
lib, rt = N.lib, C.CDLL("libcudart.so.12")
S = C.c_void_p(1)                       # legacy default stream
with ocm.LocalCluster(1, gpus=[0], gpu_bytes=256 << 20, host_bytes=64 << 20) as cl:
    cl.apply_env(0)
    ocm.init()
    n, pat = 8 << 20, 0xDEADBEEFDEADBEEF
    a = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
    bad = ocm.alloc(ocm.OCM_LOCAL_GPU, 4096)
    w = C.c_void_p(a.local_ptr)
    assert lib.ocmk_fill(w, n, C.c_uint64(pat), S) == 0 and rt.cudaDeviceSynchronize() == 0
    a.put(n)                                                      # blocking
    a.put(4096, stream=None)                                      # library stream, per-thread parameter block reused
    a.get(4096, stream=None, src_offset=8192, dest_offset=4096)
    ocm.quiet()
    assert lib.ocmk_fill(w, n, C.c_uint64(0), S) == 0 and rt.cudaDeviceSynchronize() == 0
    a.get(n)
    assert lib.ocmk_verify(w, n, C.c_uint64(pat), C.c_void_p(bad.local_ptr), S) == 0 and rt.cudaDeviceSynchronize() == 0
    host = (C.c_uint64 * 512)()
    bad.copy_out(C.addressof(host))
    assert host[0] == 0, f"{host[0]} bytes differ after put/get"
    assert a.signal() == 2, a.signal()
    h = ocm.alloc(ocm.OCM_REMOTE_HOST, 1 << 20, 1 << 20)          # host pool (pinned VMM allocation)
    C.memset(h.local_ptr, 0x5A, 1 << 20)
    h.put(1 << 20)
    C.memset(h.local_ptr, 0, 1 << 20)
    h.get(1 << 20, stream=None)
    ocm.quiet()
    assert C.string_at(h.local_ptr + (1 << 20) - 4, 4) == b"\x5a" * 4 and h.signal() == 1
    g = ocm.alloc(ocm.OCM_LOCAL_GPU, 1 << 20)
    ocm.copy(g, h, 1 << 20)                                       # host window -> GPU block
    assert lib.ocmk_verify(C.c_void_p(g.local_ptr), 1 << 20, C.c_uint64(0x5A5A5A5A5A5A5A5A), C.c_void_p(bad.local_ptr), S) == 0
    rt.cudaDeviceSynchronize()
    bad.copy_out(C.addressof(host))
    assert host[0] == 0
    for x in (a, bad, h, g):
        x.free()
    ocm.tini()
print("smoke_binding ok")
