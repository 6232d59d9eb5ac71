#!/usr/bin/env python
"""Small single-GPU run of every kernel for `compute-sanitizer --tool memcheck|racecheck|synccheck` (one tool
per gpurun call, B200_PROFILING.md).  Sizes are tiny: the sanitizers slow kernels down by orders of magnitude.
Exits non-zero if any result is wrong.
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402

import oncilla_b200 as ocm  # noqa: E402
from oncilla_b200 import ops  # noqa: E402

ok = True
n = (1 << 20) + 48
src = torch.randint(0, 256, (n,), dtype=torch.uint8, device="cuda")
for mode in (None, 1, 2):
    d1, d2 = torch.zeros_like(src), torch.zeros_like(src)
    sig = torch.zeros(1, dtype=torch.int64, device="cuda")
    ops.copy(d1, src, n, dst2=d2, signal=sig, mode=mode)
    torch.cuda.synchronize()
    ok &= torch.equal(d1, src) and torch.equal(d2, src) and int(sig.item()) == 1
ops.copy(d1[3:], src[3:], n - 3)            # head/tail handling of the 256-bit kernel
ops.copy(d1[1:], src[5:], n - 5)            # byte path
buf = torch.zeros(70001, dtype=torch.uint8, device="cuda")
ops.fill(buf[1:], 0x0102030405060708, 70000)
ok &= ops.verify(buf[1:], 0x0102030405060708, 70000) == 0
ops.checksum(buf)
# strided: bulk rows, vector rows, cast rows
s2 = torch.randint(0, 256, (256 * 8192,), dtype=torch.uint8, device="cuda")
d3 = torch.zeros(256 * 4096, dtype=torch.uint8, device="cuda")
ops.strided_copy(d3, s2, row_bytes=4096, counts=(256, 1, 1), src_strides=(8192, 0, 0), dst_strides=(4096, 0, 0))
ops.strided_copy(d3, s2, row_bytes=48, counts=(1000, 1, 1), src_strides=(64, 0, 0), dst_strides=(48, 0, 0))
x = torch.randn(64, 512, device="cuda")
q = torch.zeros(64, 256, dtype=torch.uint8, device="cuda")
sc = torch.zeros(64, 8, dtype=torch.float32, device="cuda")
ops.strided_copy(q, x, row_bytes=1024, counts=(64, 1, 1), src_strides=(2048, 0, 0), dst_strides=(256, 0, 0),
                 src_dtype=ocm.DT_F32, dst_dtype=ocm.DT_F8E4M3, scale_block=32, scales=sc)
y = torch.zeros(64, 256, device="cuda")
ops.strided_copy(y, q, row_bytes=256, counts=(64, 1, 1), src_strides=(256, 0, 0), dst_strides=(1024, 0, 0),
                 src_dtype=ocm.DT_F8E4M3, dst_dtype=ocm.DT_F32, scale_block=32, scales=sc)
torch.cuda.synchronize()
ok &= torch.allclose(y, x[:, :256], rtol=0.07, atol=1e-2)
# replicated (cluster), gups, sort
dsts = [torch.zeros(1 << 18, dtype=torch.uint8, device="cuda") for _ in range(4)]
ops.put_replicated(dsts, src, 1 << 18)
torch.cuda.synchronize()
ok &= all(torch.equal(d, src[: 1 << 18]) for d in dsts)
slabs = [torch.zeros(1 << 12, dtype=torch.int64, device="cuda") for _ in range(4)]
ops.gups(slabs, 1 << 10, 1 << 14, 1 << 15, 5)
ops.gups(slabs, 1 << 10, 1 << 14, 1 << 15, 5)
torch.cuda.synchronize()
ok &= all(int(s.abs().sum()) == 0 for s in slabs)
keys = torch.randint(0, 2**62, (50001,), dtype=torch.int64, device="cuda")
want = torch.sort(keys).values
ops.sort_u64(keys)
torch.cuda.synchronize()
ok &= torch.equal(keys, want)
print("sanitizer target:", "ok" if ok else "WRONG RESULT")
sys.exit(0 if ok else 1)
