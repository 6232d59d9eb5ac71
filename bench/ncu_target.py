#!/usr/bin/env python
"""Small single-GPU target for `ncu`: a few launches of each hot kernel on plain tensors (no daemon).

    python bench/ncu_target.py [copy|dual|vec|strided|cast|sort|gups|all]
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402

import oncilla_b200 as ocm  # noqa: E402
from oncilla_b200 import ops  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
n = 256 << 20
src = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
dst = torch.empty_like(src)
dst2 = torch.empty_like(src)
for _ in range(3):
    if what in ("v32", "all"):
        ops.copy(dst, src, n)                      # default dispatch: 256-bit vector kernel
    if what in ("v32dual", "all"):
        ops.copy(dst, src, n, dst2=dst2)
    if what in ("copy", "all"):
        ops.copy(dst, src, n, mode=1)
    if what in ("dual", "all"):
        ops.copy(dst, src, n, dst2=dst2, mode=1)
    if what in ("vec", "all"):
        ops.copy(dst, src, n, mode=2)
    if what in ("strided", "all"):
        ops.strided_copy(dst, src, row_bytes=4096, counts=(16384, 1, 1), src_strides=(8192, 0, 0), dst_strides=(4096, 0, 0))
    if what in ("cast", "all"):
        x = src.view(torch.float32)[: 32 << 20].view(8192, 4096)
        x.copy_(torch.randn(8192, 4096, device="cuda"))
        q = dst[: 8192 * 4096].view(8192, 4096)
        sc = dst2.view(torch.float32)[: 8192 * 128].view(8192, 128)
        ops.strided_copy(q, x, row_bytes=4096 * 4, counts=(8192, 1, 1), src_strides=(4096 * 4, 0, 0),
                         dst_strides=(4096, 0, 0), src_dtype=ocm.DT_F32, dst_dtype=ocm.DT_F8E4M3, scale_block=32, scales=sc)
    if what in ("sort", "all"):
        k = src.view(torch.int64)[: 1 << 22]
        ops.sort_u64(k, dst.view(torch.int64)[: 1 << 22])
    if what in ("gups", "all"):
        slabs = [dst.view(torch.int64)[i * (1 << 22):(i + 1) * (1 << 22)] for i in range(4)]
        ops.gups(slabs, 1 << 16, 1 << 24, 1 << 24, 3)
torch.cuda.synchronize()
print("ncu target done:", what)
