#!/usr/bin/env python
"""Headline benchmark: ocm_copy GB/s (GPU allocation -> remote allocation on the neighbour rank) at N GPUs.

Contract (driver): ``python bench.py --gpus N --steps K --warmup W`` (N > 1: launched under torchrun, one
rank per GPU).  One step = one ``ocm_copy`` of SIZE bytes from an OCM_LOCAL_GPU allocation into an
OCM_REMOTE_GPU pair whose far side the governor places on the neighbour rank ``(r+1) % N`` (the reference's
policy, reference src/alloc.c:107) - at N=1 that is the rank's own pool, so the copy stays inside one HBM;
at N >= 2 every byte crosses NVLink.  The kernel that runs is ``ocm_copy_v32_kernel<dual>`` (src/kernels/copy.cu):
256-bit LDG from the source, 256-bit STG to the local window and to the peer-mapped far side, followed by a
one-thread ``ocm_complete_kernel`` that publishes the completion signal (2 launches per step).  Timed on the
device with CUDA events, max over ranks, aggregate GB/s.

Besides the headline the JSON line carries
  * ``e2e``: the same copy through the blocking public API starting from pinned host memory (H2D inside), then a
    read of the 8-byte completion counter the GPU wrote to host memory;
  * ``curve``: put / get / ocm_copy GB/s from 4 KiB to 4 GiB next to cudaMemcpyAsync peer copies (copy engines,
    baseline only) in the same process on the same GPUs;
  * ``alloc_p50_us_*``: ocm_alloc latency on both allocation paths.
``--impl reference`` runs the unmodified reference built under baseline/_ref for the same metric.
"""
from __future__ import annotations

import argparse
import ctypes
import json
import os
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "baseline"))

SIZE_DEFAULT = 256 << 20     # > 126 MB L2: every iteration streams from/to HBM ("inputs larger than L2")
CURVE_SIZES = [4 << 10, 64 << 10, 1 << 20, 16 << 20, 256 << 20, 1 << 30, 4 << 30]
NVLINK_ROOFLINE_GBPS = 900.0     # BASELINE.json: NVLink 5, per GPU per direction
NVLINK_MEASURED_GBPS = 770.0     # B200_PROFILING.md: measured peer copy per direction
HBM_MEASURED_GBPS = 6544.0       # MEASURED_PEAKS.json (read + write bytes)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ocm", choices=["ocm", "reference"])
    ap.add_argument("--size", type=int, default=SIZE_DEFAULT)
    ap.add_argument("--alloc-iters", type=int, default=2000)
    ap.add_argument("--curve-max", type=int, default=CURVE_SIZES[-1], help="largest message of the size curve (0: no curve)")
    return ap.parse_args()


def dist_setup(n):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != n:
        if world == 1 and n > 1:
            raise SystemExit(f"--gpus {n} needs torchrun with --nproc-per-node {n}")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    return rank, local, world


def barrier(world):
    import torch
    import torch.distributed as dist

    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def allmax(x, world):
    import torch
    import torch.distributed as dist

    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def size_curve(ocm, ops, pair, big_src, world, cmax, flush):
    """put / get / ocm_copy and the copy-engine baseline at every size; all ranks at once, max over ranks.

    No address is reused inside one size (offsets rotate through the pair's ``cmax``-byte window and far side) and the
    L2 is flushed between measurements, so small messages are not served from cache.
    """
    import torch

    rt = ctypes.CDLL("libcudart.so.12")
    stream = torch.cuda.current_stream()
    sp = ctypes.c_void_p(int(stream.cuda_stream) or 1)
    win, far = pair.local_ptr, pair.remote_ptr
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    V = ctypes.c_void_p

    def offsets(size, iters):
        step = max(size, 1 << 20)
        span = max(cmax - size, 0) // step + 1
        return [(i % span) * step for i in range(iters)]

    def timed(fn, size, iters):
        offs = offsets(size, iters + 3)
        for o in offs[:3]:
            fn(size, o)
        flush()
        barrier(world)
        ev0.record()
        for o in offs[3:]:
            fn(size, o)
        ev1.record()
        barrier(world)
        return allmax(ev0.elapsed_time(ev1), world) / iters * 1e3      # us per message

    def ce(dst, src):
        def fn(size, o):
            if rt.cudaMemcpyAsync(V(dst + o), V(src + o), ctypes.c_size_t(size), 4, sp) != 0:    # cudaMemcpyDefault
                raise RuntimeError("cudaMemcpyAsync peer copy failed")
        return fn

    ours = {
        # ocm_copy(GPU allocation -> pair): source read once, stored to the window and to the far side
        "copy": lambda size, o: ocm.copy(pair, big_src, size, src_offset=o, dest_offset=o, src_offset_2=o, dest_offset_2=o,
                                         stream=stream),
        "put": lambda size, o: pair.put(size, src_offset=o, dest_offset=o, stream=stream),
        "get": lambda size, o: pair.get(size, src_offset=o, dest_offset=o, stream=stream),
        "ce_put": ce(far, win),
        "ce_get": ce(win, far),
    }
    rows = []
    for size in [s for s in CURVE_SIZES if s <= cmax]:
        iters = 3 if size >= 1 << 30 else 10 if size >= 64 << 20 else 40
        row = {"bytes": size}
        for name, fn in ours.items():
            try:
                us = timed(fn, size, iters)
            except RuntimeError as e:       # the baseline, not the product, failed: report and go on
                row[name + "_error"] = str(e)
                continue
            row[name + "_us"] = round(us, 2)
            row[name + "_GBps"] = round(size / us / 1e3, 2)
        rows.append(row)
    return rows


def run_ocm(args):
    import torch
    import torch.distributed as dist

    import oncilla_b200 as ocm
    from oncilla_b200 import ops
    from oncilla_b200.utils import ClockSampler
    from ref_harness import METRIC

    rank, local, world = dist_setup(args.gpus)
    S, K, W = args.size, args.steps, max(args.warmup, 3)
    cmax = args.curve_max

    # --- bring up one daemon per rank (this process starts its own) ---------------------------------
    tag = os.environ.get("MASTER_PORT", str(os.getpid()))
    ns = f"bench{tag}"
    shared = Path(f"/tmp/ocm_bench_{tag}")
    nodefile = shared / "nodefile"
    if rank == 0:
        shared.mkdir(parents=True, exist_ok=True)
        pool = max(4 * S, 2 << 30) + 3 * cmax + (256 << 20 if cmax else 0)
        ocm.write_nodefile(nodefile, world, gpus=list(range(world)), gpu_bytes=pool, host_bytes=max(2 * S, 512 << 20))
    barrier(world)
    cl = ocm.LocalCluster(world, ns=ns, nodefile=str(nodefile), ranks=[rank])
    if rank == 0:
        cl.start()
    barrier(world)
    if rank != 0:
        cl.start()
    barrier(world)
    cl.apply_env(rank)
    os.environ["OCM_DEVICE"] = str(local)
    ocm.init()
    result = None
    try:
        src = ocm.alloc(ocm.OCM_LOCAL_GPU, S)
        dst = ocm.alloc(ocm.OCM_REMOTE_GPU, S, S)
        host = ocm.alloc(ocm.OCM_LOCAL_HOST, S)       # pinned pool memory: the e2e step's input
        assert dst.remote_rank == (rank + 1) % world, (dst.remote_rank, rank, world)
        ops.fill(src.local_tensor(), 0x0123456789ABCDEF + rank)
        host.local_tensor().view(torch.int64).fill_(0x1111 + rank)
        torch.cuda.synchronize()
        stream = torch.cuda.current_stream()

        # --- device-timed: K x ocm_copy(dst <- src) -------------------------------------------------
        # clocks are sampled (nvidia-smi, 100 ms period) from the warm-up to the end of the e2e phase: the
        # device-timed region alone lasts only a few ms
        clocks = ClockSampler(local, period_ms=100)
        clocks.__enter__()
        t_warm = time.perf_counter()
        while True:                                   # at least W warm-up copies and ~0.4 s under load
            for _ in range(W):
                ocm.copy(dst, src, S, stream=stream)
            torch.cuda.synchronize()
            if time.perf_counter() - t_warm > 0.4:
                break
        barrier(world)
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        launches0 = ops.launch_count()
        barrier(world)
        ev0.record()
        for _ in range(K):
            ocm.copy(dst, src, S, stream=stream)
        ev1.record()
        barrier(world)
        launches = ops.launch_count() - launches0
        ms = allmax(ev0.elapsed_time(ev1), world)
        value = world * S * K / (ms / 1e3) / 1e9
        ok = ops.verify(dst.remote_tensor(), 0x0123456789ABCDEF + rank, S) == 0

        # --- one-sided put / get (the reference's own bandwidth test shape, test/ocm_test.c:323-425) ----
        extra = {}
        for name, fn in (("put", dst.put), ("get", dst.get)):
            for _ in range(3):
                fn(S, stream=stream)
            barrier(world)
            ev0.record()
            for _ in range(K):
                fn(S, stream=stream)
            ev1.record()
            barrier(world)
            extra[f"onesided_{name}_GBps"] = round(world * S * K / (allmax(ev0.elapsed_time(ev1), world) / 1e3) / 1e9, 2)

        # --- end to end through the public API: pinned host -> {window, far GPU}, then read the result ---
        for _ in range(2):
            ocm.copy(dst, host, S)
        sig0 = dst.signal()
        barrier(world)
        t0 = time.perf_counter()
        for i in range(K):
            ocm.copy(dst, host, S)                    # synchronous public call: H2D of S bytes is inside
            got = dst.signal()                        # 8-byte result word the GPU wrote to host memory
        t1 = time.perf_counter()
        assert got == sig0 + K
        e2e_s = allmax(t1 - t0, world)
        e2e = world * S * K / e2e_s / 1e9
        clocks.__exit__(None, None, None)
        ok = ok and ops.verify(dst.remote_tensor(), ((0x1111 + rank) & 0xFFFFFFFFFFFFFFFF), S) == 0

        # --- what the platform gives that step: K plain cudaMemcpyAsync H2D copies of the same pinned buffer into the
        # window, all ranks at once (copy engines; baseline only).  On the 8-GPU VMs of this pool the host side, not
        # PCIe, is the limit once more than two GPUs pull at the same time (profiles/h2d_ceiling_8gpu.jsonl).
        rt = ctypes.CDLL("libcudart.so.12")
        hptr, wptr = host.local_ptr, dst.local_ptr
        sp = ctypes.c_void_p(int(stream.cuda_stream) or 1)
        for _ in range(2):
            rt.cudaMemcpyAsync(ctypes.c_void_p(wptr), ctypes.c_void_p(hptr), ctypes.c_size_t(S), 1, sp)
        barrier(world)
        ev0.record()
        for _ in range(K):
            rt.cudaMemcpyAsync(ctypes.c_void_p(wptr), ctypes.c_void_p(hptr), ctypes.c_size_t(S), 1, sp)
        ev1.record()
        barrier(world)
        h2d_ceiling = world * S * K / (allmax(ev0.elapsed_time(ev1), world) / 1e3) / 1e9

        # --- GB/s versus message size, next to the copy engines (BASELINE.json metric: "GB/s vs size") ----
        curve = None
        if cmax:
            pair = ocm.alloc(ocm.OCM_REMOTE_GPU, cmax, cmax)
            big_src = ocm.alloc(ocm.OCM_LOCAL_GPU, cmax)
            assert pair.remote_rank == (rank + 1) % world
            flush_t = src.local_tensor()
            curve = size_curve(ocm, ops, pair, big_src, world, cmax, lambda: ops.fill(flush_t, 0x0123456789ABCDEF + rank))
            torch.cuda.synchronize()
            barrier(world)
            pair.free()
            big_src.free()

        # --- p50 ocm_alloc latency (fast path and daemon path) ---------------------------------------------
        def alloc_lat(flags):
            ts = []
            for _ in range(args.alloc_iters):
                t = time.perf_counter_ns()
                a = ocm.alloc(ocm.OCM_REMOTE_GPU, 0, 1 << 20, flags=flags | ocm.ALLOC_NO_WINDOW)
                ts.append(time.perf_counter_ns() - t)
                a.free()
            ts.sort()
            return ts[len(ts) // 2] / 1e3
        extra["alloc_p50_us_fast"] = round(allmax(alloc_lat(0), world), 2)
        extra["alloc_p50_us_daemon"] = round(allmax(alloc_lat(ocm.ALLOC_SLOW_PATH), world), 2)
        extra["verified"] = bool(ok)

        if rank == 0:
            per_gpu = value / world
            if world > 1:
                roof = {"frac_of_roofline": round(per_gpu / NVLINK_ROOFLINE_GBPS, 3),
                        "roofline": "NVLink 5: 900 GB/s per GPU per direction (BASELINE.json)",
                        "frac_of_measured_peer_copy": round(per_gpu / NVLINK_MEASURED_GBPS, 3),
                        "measured_peer_copy": "770 GB/s per direction (B200_PROFILING.md)"}
            else:
                roof = {"frac_of_roofline": round(per_gpu / (HBM_MEASURED_GBPS / 3), 3),
                        "roofline": "HBM: 1 read + 2 writes per byte of 6544 GB/s measured copy bandwidth "
                                    "(MEASURED_PEAKS.json); no byte crosses NVLink at 1 GPU"}
            result = {
                "metric": METRIC,
                "value": round(value, 2), "unit": "GB/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "bytes (no arithmetic)", "data": "synthetic",
                "config": {"model": "ocm_copy LOCAL_GPU->REMOTE_GPU", "message_bytes": S, "global_batch": world,
                           "seq_len": S, "parallelism": f"neighbour-put x{world}",
                           "cache": "inputs larger than L2 (no flush needed); size curve: rotating offsets + L2 flush "
                                    "between measurements",
                           "n1_note": "at 1 GPU the neighbour is the rank itself: HBM-local copy, no NVLink"},
                "clocks": clocks.summary(),
                "e2e": {"value": round(e2e, 2), "unit": "GB/s", "h2d_bytes_per_step": S, "d2h_bytes_per_step": 8,
                        "path": "ocm_copy(pinned LOCAL_HOST -> REMOTE_GPU pair), sync, then read the signal word",
                        "h2d_ceiling_GBps": round(h2d_ceiling, 2),
                        "h2d_ceiling_note": "cudaMemcpyAsync H2D of the same buffer on all ranks at once (device-timed)"},
                "gpu_launches": int(launches) * world,
                "kernels": "ocm_copy_v32_kernel<dual> + ocm_complete_kernel per step",
                **roof, "impl": "ocm", **extra,
            }
            if curve is not None:
                result["curve"] = curve
                result["curve_note"] = ("us per message and GB/s per GPU, all ranks putting/getting at once, max over "
                                        "ranks; copy = ocm_copy GPU allocation -> pair (window + far side), put/get = ocm_copy_onesided "
                                        "(our kernels), ce_* = cudaMemcpyAsync "
                                        "between the same peer-mapped buffers (copy engines, baseline only)")
        for a in (src, dst, host):
            a.free()
    finally:
        ocm.tini()
        barrier(world)
        cl.stop()
        if world > 1:
            dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(result), flush=True)


def run_reference(args):
    try:
        import ref_harness
    except Exception as e:  # pragma: no cover
        print(json.dumps({"impl": "reference", "unavailable": f"baseline harness missing: {e}"}))
        return
    ref_harness.main(args)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ocm(a)
