#!/usr/bin/env python
"""Where does the N=1 headline lose time against the bare kernel?  Dual-destination copy of 256 MiB on OCM
memory: bare kernel, kernel + device-memory signal, kernel + host-memory signal, and the public ocm_copy."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402

import oncilla_b200 as ocm  # noqa: E402
from oncilla_b200 import ops  # noqa: E402


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


n = 256 << 20
torch.cuda.set_device(0)
with ocm.LocalCluster(1, gpus=[0], gpu_bytes=2 << 30, host_bytes=64 << 20) as cl:
    cl.apply_env(0)
    ocm.init()
    src = ocm.alloc(ocm.OCM_LOCAL_GPU, n)
    dst = ocm.alloc(ocm.OCM_REMOTE_GPU, n, n)
    s, w, f = src.local_tensor(), dst.local_tensor(), dst.remote_tensor()
    dsig = torch.zeros(1, dtype=torch.int64, device="cuda")
    hsig = torch.zeros(8, dtype=torch.int64).pin_memory()
    a, b, c = (torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(3))
    st = torch.cuda.current_stream()
    rows = {
        "torch_buffers_bare": timed(lambda: ops.copy(b, a, n, dst2=c)),
        "ocm_buffers_bare": timed(lambda: ops.copy(f, s, n, dst2=w)),
        "ocm_buffers_dev_signal": timed(lambda: ops.copy(f, s, n, dst2=w, signal=dsig)),
        "ocm_buffers_host_signal": timed(lambda: ops.copy(f, s, n, dst2=w, signal=hsig)),
        "api_ocm_copy_async": timed(lambda: ocm.copy(dst, src, n, stream=st)),
        "api_put_async": timed(lambda: dst.put(n, stream=st)),
        "bare_single_dst": timed(lambda: ops.copy(f, s, n)),
    }
    for k, us in rows.items():
        print(f"{k:28s} {us:8.1f} us  {n / us / 1e3:8.1f} GB/s")
    src.free()
    dst.free()
    ocm.tini()
