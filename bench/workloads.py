#!/usr/bin/env python
"""Application-level configurations of BASELINE.json on N GPUs (run under torchrun, one rank per GPU):

  config 3  GUPS over pooled HBM: every rank updates its own table striped over ALL ranks' pools
  config 4  out-of-core sort spanning peer HBM + the pinned host tier (per rank, scaled by --sort-keys)
  config 5  strided fp8-cast copies into the neighbour's HBM (K4/K5 over NVLink)

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/workloads.py
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import oncilla_b200 as ocm  # noqa: E402
from oncilla_b200 import ops  # noqa: E402
from oncilla_b200.cluster import TorchrunCluster  # noqa: E402
from oncilla_b200.models import gups, sort  # noqa: E402


def allmax(x, world):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def allsum(x, world):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gups-words", type=int, default=1 << 28)      # 2 GiB table per rank, striped over all ranks
    ap.add_argument("--gups-updates", type=int, default=1 << 28)
    ap.add_argument("--sort-keys", type=int, default=1 << 26)       # 512 MiB of keys per rank
    ap.add_argument("--sort-chunk", type=int, default=1 << 23)
    ap.add_argument("--hbm-frac", type=float, default=0.5, help="fraction of the sort's data allowed in peer HBM; rest spills to host")
    ap.add_argument("--dsort-keys", type=int, default=1 << 24)      # 128 MiB of keys per rank, one global order
    ap.add_argument("--only", choices=["dsort", "oocsort"], default=None, help="run just this section")
    ap.add_argument("--sort-final", default="merge", choices=["merge", "sort", "both"])
    ap.add_argument("--sort-repeat", type=int, default=1, help="passes of the out-of-core sort (the first one pays the host-pool pinning)")
    ap.add_argument("--sort-parts", type=int, default=0, help="partitions per rank (0: one per run)")
    ap.add_argument("--gpu-bytes", type=str, default="8G")
    ap.add_argument("--host-bytes", type=str, default="3G")
    args = ap.parse_args()

    def size(s):
        return int(float(s[:-1]) * {"K": 1 << 10, "M": 1 << 20, "G": 1 << 30, "T": 1 << 40}[s[-1].upper()]) if s[-1].isalpha() else int(s)

    # the replicated-put section holds one 64 MiB replicated allocation per rank at once: size the NVLS region for it
    w = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("OCM_MC_BYTES", str(min(1 << 30, (64 << 20) * w + (64 << 20))))
    tc = TorchrunCluster(gpu_bytes=size(args.gpu_bytes), host_bytes=size(args.host_bytes))
    with tc as (rank, world):
        # ---- config 3 ------------------------------------------------------------------------
        tc.barrier()
        if not args.only:
            g = gups.run(total_words=args.gups_words, n_updates=args.gups_updates, iters=3)
            tc.barrier()
            worst = allmax(g["seconds"], world)
            ok = allsum(1.0 if g["verified"] else 0.0, world) == world
            tc.barrier()
            gl = gups.run(total_words=args.gups_words, n_updates=args.gups_updates, iters=3, local_only=True)
            tc.barrier()
            worst_local = allmax(gl["seconds"], world)
        if rank == 0 and not args.only:
            # Roofline of the pooled case: a remote 8-byte RED crosses NVLink as 8 bytes of payload + 32 bytes of protocol
            # (ncu nvltx__bytes_data_user / _data_protocol, profiles/nvlink_ncu_2gpu.md), so 900 GB/s per direction
            # carries at most 22.5 G remote updates/s per GPU; (world-1)/world of the updates are remote.
            per_gpu = args.gups_updates / worst / 1e9
            remote_frac = (world - 1) / world if world > 1 else 0.0
            print(json.dumps({"kind": "gups", "n_gpus": world, "table_words_per_rank": args.gups_words,
                              "updates_per_rank": args.gups_updates, "seconds_max": round(worst, 6),
                              "GUPS_aggregate": round(world * args.gups_updates / worst / 1e9, 3),
                              "GUPS_per_gpu": round(per_gpu, 3),
                              "GUPS_per_gpu_local_hbm_only": round(args.gups_updates / worst_local / 1e9, 3),
                              "nvlink_packet_bound_GUPS_per_gpu": round(900.0 / 40.0 / remote_frac, 1) if remote_frac else None,
                              "frac_of_nvlink_packet_bound": round(per_gpu * remote_frac / (900.0 / 40.0), 3) if remote_frac else None,
                              "verified": ok and bool(gl["verified"])}), flush=True)
        # ---- config 4 ------------------------------------------------------------------------
        tc.barrier()
        # half of the runs (and of the output) fit the HBM budget, the rest spills to the pinned host tier
        # (runs + output: 2 x the data; --hbm-frac of that may sit in peer HBM)
        if not args.only or args.only == "oocsort":
            from oncilla_b200.models import oocsort

            finals = ["merge", "sort"] if args.sort_final == "both" else [args.sort_final]
            for pass_no, final in enumerate(finals * args.sort_repeat):
                s = oocsort.run(keys_per_rank=args.sort_keys, chunk_keys=args.sort_chunk, parts_per_rank=args.sort_parts or None,
                                hbm_budget=int(args.hbm_frac * 2 * args.sort_keys * 8), final=final)
                tc.barrier()
                worst = allmax(s["seconds"], world)
                p1, p2 = allmax(s["phase1_s"], world), allmax(s["phase2_s"], world)
                ok = allsum(1.0 if s["ok"] else 0.0, world) == world
                if rank == 0:
                    print(json.dumps({"kind": "ooc_sort", "n_gpus": world, "keys_per_rank": args.sort_keys, "final": final,
                                      "pass": pass_no,  # pass 0 also pins the host-pool pieces it touches (once per process)
                                      "total_bytes": world * args.sort_keys * 8, "chunk_keys": args.sort_chunk,
                                      "runs_per_rank": s["n_runs"], "slices_per_partition": s["slices_per_partition"],
                                      "seconds_max": round(worst, 4), "setup_map_pools_s_rank0": round(s["setup_s"], 4),
                                      "phase1_run_formation_s": round(p1, 4),
                                      "phase2_gather_merge_s": round(p2, 4),
                                      "Mkeys_per_s_aggregate": round(world * args.sort_keys / worst / 1e6, 1),
                                      "GBps_aggregate": round(world * args.sort_keys * 8 / worst / 1e9, 2),
                                      "bytes_in_hbm_rank0": s["bytes_in_hbm"], "bytes_in_host_rank0": s["bytes_in_host"],
                                      "host_blocked_s_rank0": s["host_blocked_s"],
                                      "verified": ok}), flush=True)
        if args.only == "oocsort":
            return
        # ---- distributed sample sort: all ranks sort one data set, keys exchanged by one-sided puts into
        # shared blocks (ocm_share / ocm_attach) with completion signals ------------------------------------
        from oncilla_b200.models import dsort

        for it in range(2):                                   # first pass warms allocations and NCCL
            d = dsort.run(keys_per_rank=args.dsort_keys)
        if rank == 0:
            print(json.dumps({"kind": "distributed_sort", "n_gpus": world, "keys_per_rank": args.dsort_keys,
                              "seconds_max": round(d["seconds"], 5), "local_sort_s": round(d["local_sort_s"], 5),
                              "exchange_s": round(d["exchange_s"], 5), "final_sort_s": round(d["final_sort_s"], 5),
                              "Mkeys_per_s_aggregate": round(d["mkeys_per_s"], 1), "verified": d["ok"]}), flush=True)
        if args.only == "dsort":
            return
        # ---- config 5 (copy part): strided fp32 -> block-scaled fp8 put into the neighbour's HBM ----------
        rows, elems, pitch = 16384, 4096, 8192
        src = ocm.alloc(ocm.OCM_LOCAL_GPU, rows * pitch * 4)
        src.local_tensor(torch.float32).normal_()
        payload = rows * elems
        dst = ocm.alloc(ocm.OCM_REMOTE_GPU, 0, payload + rows * (elems // 32) * 4, flags=ocm.ALLOC_NO_WINDOW)
        st = torch.cuda.current_stream()

        def step():
            ocm.copy_strided(dst, src, row_bytes=elems * 4, counts=(rows, 1, 1), src_strides=(pitch * 4, 0, 0),
                             dst_strides=(elems, 0, 0), src_dtype=ocm.DT_F32, dst_dtype=ocm.DT_F8E4M3, scale_block=32,
                             scale_offset=payload, stream=st)

        for _ in range(3):
            step()
        tc.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            step()
        e1.record()
        tc.barrier()
        sec = allmax(e0.elapsed_time(e1) / 1e3 / 10, world)
        if rank == 0:
            print(json.dumps({"kind": "strided_fp8_put", "n_gpus": world, "rows": rows, "elems": elems,
                              "src_bytes": rows * elems * 4, "wire_bytes": payload + rows * (elems // 32) * 4,
                              "us": round(sec * 1e6, 1),
                              "src_GBps_aggregate": round(world * rows * elems * 4 / sec / 1e9, 1),
                              "wire_GBps_per_gpu": round((payload + rows * (elems // 32) * 4) / sec / 1e9, 1)}), flush=True)
        src.free()
        dst.free()
        # ---- replicated put (K6): one source -> every GPU; rank 0 alone, then all ranks at once ------------
        from oncilla_b200 import api as _api

        nrep = 64 << 20
        ready = _api.multicast_ready(wait_s=5.0) if world > 1 else False
        rep = ocm.alloc(ocm.OCM_REPLICATED_GPU, nrep, nrep)
        slabs = [rep.slab_tensor(r) for r in range(world)]
        win = rep.local_tensor()
        ops.fill(win, 0x1234)
        torch.cuda.synchronize()

        def timeit(fn, everyone):
            tc.barrier()
            if everyone or rank == 0:
                for _ in range(3):
                    fn()
                torch.cuda.synchronize()
            tc.barrier()
            e0.record()
            if everyone or rank == 0:
                for _ in range(10):
                    fn()
            e1.record()
            torch.cuda.synchronize()
            tc.barrier()
            return allmax(e0.elapsed_time(e1) / 1e3 / 10, world)

        rows_out = {}
        for everyone in (False, True):
            tag = "all_ranks" if everyone else "rank0_only"
            rows_out[tag] = {
                "api_put_us": round(timeit(lambda: rep.put(nrep, stream=st), everyone) * 1e6, 1),
                "cluster_kernel_us": round(timeit(lambda: ops.put_replicated(slabs[:8], win, nrep), everyone) * 1e6, 1),
                "unicast_loop_us": round(timeit(lambda: [ops.copy(s, win, nrep) for s in slabs], everyone) * 1e6, 1),
            }
        if rank == 0:
            print(json.dumps({"kind": "replicated_put", "n_gpus": world, "bytes": nrep, "nvls": bool(rep.is_multicast),
                              "multicast_ready": ready, **rows_out}), flush=True)
        rep.free()


if __name__ == "__main__":
    main()
