"""SURVEY Appendix B, item by item: the reference defects that must not be reproduced (host tiers, no GPU needed).

Numbers in the assertion messages are the Appendix-B item numbers; `ref:` cites jyoung3131/oncilla.
"""
import json

from conftest import run_py

SNIPPET = r"""
import ctypes as C, json, torch
import oncilla_b200 as ocm
from oncilla_b200 import _native as N
lib = N.lib
ocm.init()
res = {}
loc = ocm.alloc(ocm.OCM_LOCAL_HOST, 1 << 20)
rem = ocm.alloc(ocm.OCM_REMOTE_HOST, 1 << 20, 48 << 20)
rdma = ocm.alloc(ocm.OCM_REMOTE_RDMA, 4096, 4096)
# 1. ocm_is_remote (ref:src/lib.c:461 answers true only for OCM_LOCAL_GPU)
res["is_remote"] = [loc.is_remote, rem.is_remote, rdma.is_remote]
# 2. ocm_free(NULL) / double free (ref:src/lib.c:357-359 dereferences first)
res["free_null"] = lib.ocm_free(None)
h = lib.ocm_alloc_ex(C.byref(N.ocm_alloc_params_ex()))            # kind 0: refused, NULL handle, no crash
res["alloc_kind0"] = h
tmp = ocm.alloc(ocm.OCM_LOCAL_HOST, 4096); raw = tmp.handle
res["free_once"] = lib.ocm_free(raw); tmp._h = None
res["free_twice"] = lib.ocm_free(raw)
# 3. OCM_LOCAL_HOST allocations are tracked (ref:src/lib.c:222-229 never lists them)
res["live_counts_local"] = ocm.stats()["allocs_live"]
# 4. ocm_copy must not change the caller's op_flag (ref:src/lib.c:511-514 flips it)
p = N.ocm_params(); p.bytes = 4096; p.op_flag = 0
rc = lib.ocm_copy(loc.handle, rem.handle, C.byref(p))
res["op_flag_after_read"] = [rc, p.op_flag]
# 5. ocm_copy_in / ocm_copy_out exist (ref:src/lib.c:491-499 return -1)
buf = (C.c_uint8 * (1 << 20))(*([7] * 16)); out = (C.c_uint8 * (1 << 20))()
res["copy_in_out"] = [lib.ocm_copy_in(loc.handle, buf), lib.ocm_copy_out(out, loc.handle), out[0], out[15], out[16]]
# 10/11. transfers beyond 16 MiB and beyond 4 GiB-safe arithmetic move every byte at the right offsets
#        (ref:src/extoll.c:71-106 rewrites the first 16 MiB; ref:src/rdma.c:62 truncates to 32 bits)
big = ocm.alloc(ocm.OCM_REMOTE_HOST, 40 << 20, 48 << 20)
w = big.local_tensor(); w[:] = 1; w[(17 << 20):(17 << 20) + 3] = torch.tensor([9, 8, 7], dtype=torch.uint8); w[-1] = 5
big.put(40 << 20, dest_offset=8 << 20)
far = big.remote_tensor()
res["big_put"] = [int(far[(8 << 20) - 1]), int(far[8 << 20]), far[(25 << 20):(25 << 20) + 3].tolist(), int(far[(48 << 20) - 1])]
# 17. capacity accounting and release bookkeeping on rank 0 (ref:src/alloc.c:87-92, src/mem.c:220-229: none)
used0 = ocm.stats(rank=1)["pool_host_used"]
try:
    ocm.alloc(ocm.OCM_REMOTE_HOST, 0, 1 << 40, flags=ocm.ALLOC_NO_WINDOW); res["overcommit"] = "accepted"
except ocm.OcmError as e:
    res["overcommit"] = e.err
big.free(); rem.free()
import time; time.sleep(0.3)
res["used_drop"] = used0 - ocm.stats(rank=1)["pool_host_used"]
for a in (loc, rdma): a.free()
res["live_end"] = ocm.stats()["allocs_live"]
ocm.tini()
print("RESULT " + json.dumps(res))
"""


def test_appendix_b_defects_are_not_reproduced():
    import oncilla_b200 as ocm

    with ocm.LocalCluster(2, host_bytes=192 << 20, extra_env={"OCM_NO_GPU": "1"}, ns=f"defects{__import__('os').getpid()}") as cl:
        r = run_py(cl, 0, SNIPPET)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    res = json.loads(r.stdout.split("RESULT ", 1)[1])
    assert res["is_remote"] == [False, True, True], "1"
    assert res["free_null"] == -1 and not res["alloc_kind0"], "2"
    assert res["free_once"] == 0 and res["free_twice"] == -1, "2"
    assert res["live_counts_local"] >= 3, "3"
    assert res["op_flag_after_read"] == [0, 0], "4"
    assert res["copy_in_out"] == [0, 0, 7, 7, 0], "5"
    assert res["big_put"] == [0, 1, [9, 8, 7], 5], "10/11"
    assert res["overcommit"] == 2, "17 (OCM_ENOMEM)"
    assert res["used_drop"] >= (48 << 20) * 2 and res["live_end"] == 0, "17"
