#!/usr/bin/env python
"""Topology probe for the pinned-host tier: NUMA nodes, the node each GPU hangs off, whether mbind/set_mempolicy are
permitted in this container, and the H2D copy rate from memory first-touched on every node (the ceiling of the e2e step).
Prints JSON lines; run on the GPU box (`python bench/topo_probe.py [gpu ...]`)."""
import ctypes, glob, json, os, subprocess, sys, time

def sh(cmd):
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=30).stdout.strip()
    except Exception as e:
        return f"ERR {e}"

def cpulist(s):
    out = []
    for part in s.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        out += list(range(int(a), int(b or a) + 1))
    return out

nodes = {}
for p in sorted(glob.glob("/sys/devices/system/node/node[0-9]*")):
    n = int(p.rsplit("node", 1)[1])
    nodes[n] = cpulist(open(p + "/cpulist").read())
    mem = sh(f"grep -E 'MemTotal|MemFree' {p}/meminfo")
    print(json.dumps({"node": n, "cpus": len(nodes[n]), "first_cpu": nodes[n][:1], "mem": mem.split("\n")}))
print(json.dumps({"affinity_now": len(os.sched_getaffinity(0)), "nproc": os.cpu_count()}))
gpus = []
for d in sorted(glob.glob("/sys/bus/pci/devices/*")):
    try:
        if open(d + "/vendor").read().strip() == "0x10de" and open(d + "/class").read().startswith("0x0302"):
            gpus.append({"bdf": os.path.basename(d), "numa_node": int(open(d + "/numa_node").read()),
                         "local_cpulist": open(d + "/local_cpulist").read().strip()})
    except OSError:
        pass
print(json.dumps({"pci_gpus": gpus}))
print(sh("nvidia-smi topo -m | head -30"))
print(sh("nvidia-smi --query-gpu=index,pci.bus_id,pcie.link.gen.current,pcie.link.width.current --format=csv"))

libc = ctypes.CDLL(None, use_errno=True)
SYS_mbind, SYS_set_mempolicy, SYS_get_mempolicy = 237, 238, 239
mask = ctypes.c_ulong(1)
r = libc.syscall(SYS_set_mempolicy, 1, ctypes.byref(mask), 65)   # MPOL_PREFERRED node 0
print(json.dumps({"set_mempolicy_rc": r, "errno": ctypes.get_errno()}))
libc.syscall(SYS_set_mempolicy, 0, None, 0)
buf = ctypes.create_string_buffer(1 << 20)
addr = (ctypes.addressof(buf) + 4095) & ~4095
r = libc.syscall(SYS_mbind, ctypes.c_void_p(addr), 1 << 16, 2, ctypes.byref(mask), 65, 0)   # MPOL_BIND node 0
print(json.dumps({"mbind_rc": r, "errno": ctypes.get_errno()}))

import torch
if not torch.cuda.is_available():
    sys.exit(0)
print(sh("free -g | head -3; df -h /dev/shm | tail -1; nproc"))
from pathlib import Path
REPO = Path(__file__).resolve().parent.parent
rt = ctypes.CDLL("libcudart.so.12")
ocm = ctypes.CDLL(str(REPO / "build" / "lib" / "libocm.so"))
ocm.ocmk_copy.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_size_t, ctypes.c_void_p, ctypes.c_void_p]
V = ctypes.c_void_p
def memcpy_async(dst, src, n, kind, stream):
    assert rt.cudaMemcpyAsync(V(dst), V(src), ctypes.c_size_t(n), kind, V(stream)) == 0
devs = [int(a) for a in sys.argv[1:]] or list(range(torch.cuda.device_count()))
S = 512 << 20
R = 6
all_cpus = os.sched_getaffinity(0)

def host_buffers(kind, n):
    """n pinned buffers of S bytes: 'alloc' = cudaHostAlloc, 'shm' = tmpfs mapping + cudaHostRegister (what the daemons do)."""
    out = []
    for i in range(n):
        if kind == "alloc":
            p = V()
            assert rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(S), 1 | 2) == 0      # portable | mapped
            ctypes.memset(p, 1, S)
            out.append(p.value)
        else:
            import mmap
            fd = os.open(f"/dev/shm/ocm_probe_{os.getpid()}_{i}", os.O_RDWR | os.O_CREAT, 0o600)
            os.ftruncate(fd, S)
            m = mmap.mmap(fd, S)
            os.close(fd)
            os.unlink(f"/dev/shm/ocm_probe_{os.getpid()}_{i}")
            a = ctypes.addressof(ctypes.c_char.from_buffer(m))
            ctypes.memset(a, 1, S)
            assert rt.cudaHostRegister(V(a), ctypes.c_size_t(S), 1 | 2) == 0
            out.append(a)
            host_buffers.keep.append(m)
    return out
host_buffers.keep = []

dbuf, streams = {}, {}
for d in devs:
    torch.cuda.set_device(d)
    dbuf[d] = torch.empty(S, dtype=torch.uint8, device=f"cuda:{d}")
    streams[d] = torch.cuda.Stream(device=d)

def run(kind_name, bufs, group, how):
    """All GPUs of `group` copy at once; returns per-GPU GB/s (CUDA events per device)."""
    evs = {}
    for rep in range(2):     # first pass = warm-up
        for d in group:
            torch.cuda.set_device(d)
            st = streams[d]
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(R if rep else 1):
                if how == "h2d":
                    memcpy_async(dbuf[d].data_ptr(), bufs[d], S, 1, st.cuda_stream)
                elif how == "d2h":
                    memcpy_async(bufs[d], dbuf[d].data_ptr(), S, 2, st.cuda_stream)
                else:        # our copy kernel reading the pinned buffer directly (zero-copy)
                    assert ocm.ocmk_copy(dbuf[d].data_ptr(), None, bufs[d], S, None, st.cuda_stream) == 0
            e1.record(st)
            evs[d] = (e0, e1)
        for d in group:
            torch.cuda.synchronize(d)
    per = {d: round(R * S / (evs[d][0].elapsed_time(evs[d][1]) / 1e3) / 1e9, 1) for d in group}
    print(json.dumps({"mem": kind_name, "how": how, "n_gpus": len(group), "per_gpu_GBps": per,
                      "sum": round(sum(per.values()), 1)}), flush=True)

for kind in ("alloc", "shm"):
    bufs = dict(zip(devs, host_buffers(kind, len(devs))))
    n = 1
    while n <= len(devs):
        for how in ("h2d", "kernel", "d2h"):
            run(kind, bufs, devs[:n], how)
        n *= 2
