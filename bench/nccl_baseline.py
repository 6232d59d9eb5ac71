#!/usr/bin/env python
"""Library baseline B0: NCCL send/recv bandwidth vs message size between neighbour ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/nccl_baseline.py

Every rank sends to (r+1) % N and receives from (r-1) % N at the same time - the same traffic pattern as
bench.py.  Device-timed with CUDA events, max over ranks, aggregate GB/s.  Baseline only: never on the
product path.
"""
import json
import os

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    big = 1 << 30
    sbuf = torch.ones(big, dtype=torch.uint8, device="cuda")
    rbuf = torch.empty(big, dtype=torch.uint8, device="cuda")
    for n in [4 << 10 << (2 * i) for i in range(10)]:
        iters = 50 if n <= (1 << 20) else 10 if n <= (64 << 20) else 4

        def step():
            ops = [dist.P2POp(dist.isend, sbuf[:n], nxt), dist.P2POp(dist.irecv, rbuf[:n], prv)]
            for w in dist.batch_isend_irecv(ops):
                w.wait()

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            step()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 1e3 / iters], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"kind": "nccl_sendrecv", "n_gpus": world, "bytes": n, "us": round(t.item() * 1e6, 2),
                              "GBps_aggregate": round(world * n / t.item() / 1e9, 2),
                              "GBps_per_gpu": round(n / t.item() / 1e9, 2)}), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
