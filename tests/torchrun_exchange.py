"""Helper launched by tests/test_cluster_cpu.py under torchrun (one process per rank, no GPU needed):

every rank brings up its daemon through ``TorchrunCluster``, allocates a receive block in its own pool, shares it,
attaches to everybody else's, and writes its own slot of every peer's block with one-sided copies - the all-to-all
pattern of ``models/dsort.py`` on the pinned-host tier.  Each rank then waits for its block's signal to count every
sender and checks what arrived.  ``torch.distributed`` (gloo here) only carries the tokens.
"""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import oncilla_b200 as ocm  # noqa: E402
from oncilla_b200.cluster import TorchrunCluster  # noqa: E402

SLOT = 64 << 10


def main():
    with TorchrunCluster(host_bytes=64 << 20) as (rank, world):
        recv = ocm.alloc(ocm.OCM_LOCAL_HOST, world * SLOT)
        tokens = [None] * world
        dist.all_gather_object(tokens, recv.share())
        peers = [ocm.attach(t) for t in tokens]
        src = ocm.alloc(ocm.OCM_LOCAL_HOST, SLOT)
        src.local_tensor()[:] = 100 + rank
        for k in range(world):
            dst = (rank + 1 + k) % world
            ocm.copy(peers[dst], src, SLOT, dest_offset=rank * SLOT)       # my slot in dst's block
        recv.wait_signal(world, timeout_ms=20000)                           # one copy from every rank, myself included
        t = recv.local_tensor()
        ok = all(int(t[r * SLOT]) == 100 + r and int(t[(r + 1) * SLOT - 1]) == 100 + r for r in range(world))
        far = ocm.alloc(ocm.OCM_REMOTE_HOST, 4096, 4096)                    # neighbour placement across the job's daemons
        oks = [None] * world
        dist.all_gather_object(oks, [ok, far.remote_rank, recv.signal()])
        for p in peers:
            p.free()
        dist.barrier()
        for a in (far, src, recv):
            a.free()
        if rank == 0:
            print("RESULT " + json.dumps(oks), flush=True)


if __name__ == "__main__":
    main()
