"""`bench.py --impl reference`: run the UNMODIFIED reference (built into baseline/_ref by baseline/Makefile)
through its own public API for the same metric as bench.py's own arm.

Layout per torchrun rank r (one per GPU): a private UTS+IPC namespace (`unshare --uts --ipc`) with hostname
``ocmref<r>`` - the reference finds its rank by hostname and has ONE global daemon queue per IPC namespace
(reference src/nodefile.c:92-103, inc/pmsg.h:24), so several daemons on one host need this - running the
reference daemon plus `ref_bench`, a timing harness that only calls the reference's ocm_* functions.  The
nine ib_* symbols come from the loop-back shim (baseline/ib_shim.c).  With one rank the reference itself
downgrades the remote request to local host memory (reference src/alloc.c:82-83): its N=1 number is a
pageable cudaMemcpy D2H; with N >= 2 it is cudaMemcpy D2H into the bounce buffer + ib_write to the
neighbour daemon (reference src/lib.c:643-649).
"""
from __future__ import annotations

import json
import os
import shutil
import socket
import subprocess
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
# the ONE metric string both arms of bench.py print (bench.py imports it from here)
METRIC = "ocm_copy GB/s (GPU alloc -> remote alloc on neighbour rank), aggregate over ranks"
REF = HERE / "_ref"


def _free_ports(n):
    """n free TCP ports BELOW the ephemeral range, so no outgoing connection can grab one before the daemon binds."""
    import random

    ports = []
    rng = random.Random(os.getpid() ^ int(time.time() * 1000))
    while len(ports) < n:
        p = rng.randrange(10000, 30000)
        if p in ports or 2400 <= p <= 2600:
            continue
        s = socket.socket()
        try:
            s.bind(("0.0.0.0", p))
            ports.append(p)
        except OSError:
            pass
        finally:
            s.close()
    return ports


def _have_cuda_device() -> bool:
    """True when the CUDA driver reports at least one device (asked directly: no torch import, no /dev guessing)."""
    import ctypes

    try:
        cu = ctypes.CDLL("libcuda.so.1")
        n = ctypes.c_int(0)
        return cu.cuInit(0) == 0 and cu.cuDeviceGetCount(ctypes.byref(n)) == 0 and n.value > 0
    except OSError:
        return False


def unavailable(why: str) -> None:
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)


def main(args) -> None:
    if not (REF / "bin" / "oncillamem").exists() or not (REF / "bin" / "ref_bench").exists():
        return unavailable("baseline/_ref not built (make -C baseline REF=/root/reference); pip install of the "
                           "reference is impossible: it has neither setup.py nor pyproject.toml")
    if shutil.which("unshare") is None:
        return unavailable("unshare(1) missing: cannot give each reference daemon its own hostname/IPC namespace")
    if not _have_cuda_device():
        return unavailable("no CUDA device on this machine: the reference ignores every cudaMalloc/cudaMemcpy error "
                           "except out-of-memory (src/lib.c:240-249,640), so its GPU copy would 'run' and time nothing")
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        return unavailable(f"--gpus {args.gpus} needs torchrun with {args.gpus} processes (WORLD_SIZE={world})")
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo")

    def barrier():
        if dist:
            dist.barrier()

    def bcast(obj):
        if not dist:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=0)
        return box[0]

    S, K, W = args.size, args.steps, max(args.warmup, 3)
    gathered = None
    for attempt in range(2):       # the reference survives an 8-rank run about two times out of three: retry once
        gathered = _attempt(args, rank, local, world, dist, barrier, attempt)
        if not any("error" in g for g in gathered):
            break
    _report(rank, world, gathered, S, K, W)
    if dist:
        dist.destroy_process_group()


def _attempt(args, rank, local, world, dist, barrier, attempt):
    tag = os.environ.get("MASTER_PORT", str(os.getpid())) + f"_{attempt}"
    shared = Path(f"/tmp/ocm_ref_{tag}")
    if rank == 0:
        shutil.rmtree(shared, ignore_errors=True)
        shared.mkdir(parents=True)
        ports = _free_ports(world)
        lines = ["#rank dns ethernet_ip ocm_port rdmacm_port"]
        lines += [f"{r} ocmref{r} 127.0.0.1 {ports[r]} 67890" for r in range(world)]
        (shared / "nodefile").write_text("\n".join(lines) + "\n")
    barrier()
    nodefile, go, log = shared / "nodefile", shared / "go", shared / f"daemon{rank}.log"
    S, K, W = args.size, args.steps, max(args.warmup, 3)
    script = (f"hostname ocmref{rank} || exit 97; {REF}/bin/oncillamem {nodefile} > {log} 2>&1 & D=$!; "
              f"for i in $(seq 1 3000); do [ -e {go} ] && break; sleep 0.02; done; "
              # alloc-latency loop on rank 0 only: every placement adds a dangling node to the reference's
              # rank-0 list, the fewer the better for its survival at 8 ranks
              f"{REF}/bin/ref_bench {S} {K} {W} {min(args.alloc_iters, 20) if rank == 0 else 0}; RC=$?; "
              f"kill -INT $D 2>/dev/null; sleep 0.2; kill -KILL $D 2>/dev/null; exit $RC")
    env = dict(os.environ)
    env["CUDA_VISIBLE_DEVICES"] = str(local)     # the reference always uses device 0
    env["REF_SYNC_DIR"], env["REF_RANK"], env["REF_WORLD"] = str(shared), str(rank), str(world)
    env.pop("OCM_VERBOSE", None)

    def launch():
        return subprocess.Popen(["unshare", "--uts", "--ipc", "bash", "-c", script], env=env, stdout=subprocess.PIPE,
                                stderr=subprocess.PIPE, text=True, start_new_session=True)

    proc = None
    if rank == 0:
        proc = launch()            # master first ("start the master first", reference src/mem.c:472)
        time.sleep(0.5)
    barrier()
    if rank != 0:
        proc = launch()
    barrier()
    time.sleep(1.0)                # MSG_ADD_NODE from every rank has reached rank 0
    barrier()
    if rank == 0:
        go.touch()
    try:
        out, err = proc.communicate(timeout=int(os.environ.get("OCM_REF_TIMEOUT_S", "75")))
    except subprocess.TimeoutExpired:
        # the reference blocks forever on a lost peer: never let it hang the caller
        try:
            os.killpg(os.getpgid(proc.pid), 9)
        except OSError:
            proc.kill()
        out, err = proc.communicate()
    res = None
    for line in out.splitlines():
        if line.startswith("{"):
            try:
                res = json.loads(line)
            except json.JSONDecodeError:
                pass
    mine = res if (res and "error" not in res and proc.returncode == 0) else {
        "error": (res or {}).get("error") or f"rc={proc.returncode} out={out[-300:]!r} err={err[-300:]!r}"}
    gathered = [mine]
    if dist:
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
    barrier()
    return gathered


def _report(rank, world, gathered, S, K, W):
    if rank == 0:
        bad = [g for g in gathered if "error" in g]
        if bad:
            unavailable(f"reference run failed at {world} rank(s) (its daemons hang or abort on their own list/"
                        "counter races once several ranks allocate, SURVEY App. B #7/#12): " + str(bad[0]["error"])[:200])
        else:
            copy_s = max(g["copy_s"] for g in gathered)
            e2e_s = max(g["e2e_s"] for g in gathered)
            out = {
                "impl": "reference",
                "metric": METRIC,
                "value": round(world * S * K / copy_s / 1e9, 3), "unit": "GB/s", "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": round(copy_s / K * 1e3, 4), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bytes (no arithmetic)", "data": "synthetic",
                "config": {"model": "ocm_copy LOCAL_GPU->REMOTE_RDMA (ib loop-back shim)", "message_bytes": S,
                           "global_batch": world, "seq_len": S, "parallelism": f"neighbour-put x{world}",
                           "granted_kind": gathered[0]["granted_kind"], "timing": "host clock around blocking calls"},
                # like bench.py's own arm: host allocation -> local GPU -> "remote"; the reference's GPU -> remote hop
                # is a pageable cudaMemcpy D2H of the whole message into its bounce buffer (reference src/lib.c:645)
                "e2e": {"value": round(world * S * K / e2e_s / 1e9, 3), "unit": "GB/s", "h2d_bytes_per_step": S,
                        "d2h_bytes_per_step": S,
                        "path": "ocm_copy(LOCAL_HOST -> LOCAL_GPU) then ocm_copy(LOCAL_GPU -> remote), blocking calls"},
                "e2e_host_to_remote_GBps": round(world * S * K / max(g["h2r_s"] for g in gathered) / 1e9, 3),
                "gpu_launches": 0,
                "alloc_p50_us": max(g["alloc_p50_us"] for g in gathered),
            }
            if gathered[0].get("put_s"):
                out["onesided_put_GBps"] = round(world * S * K / max(g["put_s"] for g in gathered) / 1e9, 3)
                out["onesided_get_GBps"] = round(world * S * K / max(g["get_s"] for g in gathered) / 1e9, 3)
            print(json.dumps(out), flush=True)
