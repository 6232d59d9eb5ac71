"""Launch and manage a set of ``oncillamem`` daemons on this host.

The reference needs one real machine per rank (rank = hostname match, one global
daemon queue per host: reference src/nodefile.c:92-103, inc/pmsg.h:24) and refers
to a ``bin/launch`` script that does not exist (reference README:42-43).  Here N
ranks (one per GPU) run on one box: ``OCM_RANK`` picks the rank and ``OCM_NS``
isolates the IPC names of concurrent clusters.
"""
from __future__ import annotations

import os
import signal
import socket
import subprocess
import tempfile
import time
from pathlib import Path

from . import _native


def _free_ports(n: int) -> list[int]:
    """n free TCP ports below the ephemeral range (an outgoing connection cannot grab one before the daemon binds)."""
    import random

    ports: list[int] = []
    rng = random.Random(os.getpid() ^ int(time.time() * 1000))
    while len(ports) < n:
        p = rng.randrange(10000, 30000)
        if p in ports:
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


def write_nodefile(path: str | Path, n_ranks: int, ports: list[int] | None = None, gpus: list[int] | None = None,
                   gpu_bytes: int | None = None, host_bytes: int | None = None, hostname: str = "localhost") -> list[int]:
    ports = ports or _free_ports(n_ranks)
    lines = ["#rank dns ethernet_ip ocm_port rdmacm_port [gpu=<i> gpu_bytes=<n> host_bytes=<n>]"]
    for r in range(n_ranks):
        opts = []
        if gpus is not None:
            opts.append(f"gpu={gpus[r]}")
        if gpu_bytes is not None:
            opts.append(f"gpu_bytes={int(gpu_bytes)}")
        if host_bytes is not None:
            opts.append(f"host_bytes={int(host_bytes)}")
        lines.append(f"{r} {hostname} 127.0.0.1 {ports[r]} 0 " + " ".join(opts))
    Path(path).write_text("\n".join(lines) + "\n")
    return ports


class TorchrunCluster:
    """Inside a torchrun job (one process per GPU): bring up one daemon per rank and attach this process.

    Rank 0 writes the shared nodefile, every rank starts ITS daemon (master first), then ``ocm.init()``.
    ``with TorchrunCluster(gpu_bytes=...) as (rank, world): ...``
    """

    def __init__(self, gpu_bytes: int = 4 << 30, host_bytes: int = 1 << 30, backend: str = "nccl"):
        self.gpu_bytes, self.host_bytes, self.backend = gpu_bytes, host_bytes, backend
        self.cl = None

    def barrier(self):
        import torch
        import torch.distributed as dist

        if self.world > 1:
            dist.barrier()
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def __enter__(self):
        import torch
        import torch.distributed as dist

        from . import api

        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", str(self.rank)))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.gpu = torch.cuda.is_available()       # without a GPU: host tiers only, gloo for the control plane
        if self.gpu:
            torch.cuda.set_device(self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            if self.gpu:
                dist.init_process_group(self.backend, device_id=torch.device(f"cuda:{self.local}"))
            else:
                dist.init_process_group("gloo")
        tag = os.environ.get("MASTER_PORT", str(os.getpid()))
        shared = Path(f"/tmp/ocm_job_{tag}")
        nodefile = shared / "nodefile"
        if self.rank == 0:
            shared.mkdir(parents=True, exist_ok=True)
            write_nodefile(nodefile, self.world, gpus=list(range(self.world)) if self.gpu else None,
                           gpu_bytes=self.gpu_bytes if self.gpu else None, host_bytes=self.host_bytes)
        self.barrier()
        self.cl = LocalCluster(self.world, ns=f"job{tag}", nodefile=str(nodefile), ranks=[self.rank],
                               extra_env=None if self.gpu else {"OCM_NO_GPU": "1"})
        # creating, zeroing and (in the shared-memory form) pinning the pools takes time in proportion to their size
        wait = 60.0 + 4.0 * (self.host_bytes >> 30) + 0.5 * (self.gpu_bytes >> 30)
        if self.rank == 0:
            self.cl.start(timeout=wait)
        self.barrier()
        if self.rank != 0:
            self.cl.start(timeout=wait)
        self.barrier()
        self.cl.apply_env(self.rank)
        if self.gpu:
            os.environ["OCM_DEVICE"] = str(self.local)
        else:
            os.environ["OCM_NO_GPU"] = "1"
        api.init()
        return self.rank, self.world

    def __exit__(self, *exc):
        from . import api

        if exc and exc[0] is not None:
            # A rank that failed must not wait for the others in a barrier they will never reach (that hides
            # the error behind a hang): report, stop our daemon and leave - torchrun then tears the job down.
            import sys
            import traceback

            traceback.print_exception(*exc)
            sys.stderr.flush()
            try:
                api.tini()
            finally:
                self.cl.stop()
                os._exit(1)
        api.tini()
        self.barrier()
        self.cl.stop()
        if self.world > 1:
            import torch.distributed as dist

            dist.destroy_process_group()


class LocalCluster:
    """N daemons on this host.  Use as a context manager.

    ``ranks`` restricts which ranks THIS process starts (a torchrun rank starts only its own daemon
    and shares the nodefile with its peers through ``nodefile=``).
    """

    def __init__(self, n_ranks: int = 1, *, gpus: list[int] | None | str = "auto", gpu_bytes: int | None = None,
                 host_bytes: int | None = 64 << 20, ns: str | None = None, nodefile: str | None = None,
                 ranks: list[int] | None = None, log_level: str | None = None, extra_env: dict | None = None,
                 ports: list[int] | None = None):
        self.n_ranks = n_ranks
        self.ns = ns if ns is not None else f"c{os.getpid()}"
        self.ranks = list(range(n_ranks)) if ranks is None else list(ranks)
        self.procs: dict[int, subprocess.Popen] = {}
        self._tmp = None
        self.log_level = log_level or os.environ.get("OCM_LOG", "warn")
        self.extra_env = dict(extra_env or {})
        if gpus == "auto":
            gpus = None  # daemon picks rank % device_count (or none without a GPU)
        self.gpus = gpus
        if nodefile is None:
            self._tmp = tempfile.TemporaryDirectory(prefix="ocm_")
            nodefile = os.path.join(self._tmp.name, "nodefile")
            write_nodefile(nodefile, n_ranks, ports=ports, gpus=gpus, gpu_bytes=gpu_bytes, host_bytes=host_bytes)
        self.nodefile = nodefile
        self.logs: dict[int, str] = {}

    def env(self, rank: int = 0) -> dict:
        e = dict(os.environ)
        e.update(self.extra_env)
        e["OCM_NS"] = self.ns
        e["OCM_RANK"] = str(rank)
        e["OCM_LOG"] = self.log_level
        e["OCM_DIE_WITH_PARENT"] = "1"      # read by the daemons only: they exit when this harness dies
        return e

    def apply_env(self, rank: int = 0) -> None:
        """Point THIS process at ``rank``'s daemon (call before ``ocm.init()``)."""
        os.environ["OCM_NS"] = self.ns
        os.environ["OCM_RANK"] = str(rank)

    def start(self, timeout: float = 60.0) -> "LocalCluster":
        logdir = self._tmp.name if self._tmp else tempfile.mkdtemp(prefix="ocm_logs_")
        for r in sorted(self.ranks):   # rank 0 (the master) first
            log = os.path.join(logdir, f"daemon{r}.log")
            self.logs[r] = log
            with open(log, "w") as f:
                self.procs[r] = subprocess.Popen([str(_native.DAEMON_PATH), self.nodefile], env=self.env(r), stdout=f,
                                                 stderr=subprocess.STDOUT, start_new_session=True)
            self._wait_ready(r, timeout)
        return self

    def _wait_ready(self, r: int, timeout: float) -> None:
        t0 = time.time()
        while time.time() - t0 < timeout:
            if self.procs[r].poll() is not None:
                raise RuntimeError(f"daemon {r} exited with {self.procs[r].returncode}:\n{self.read_log(r)}")
            if "oncillamem ready" in self.read_log(r):
                return
            time.sleep(0.01)
        raise TimeoutError(f"daemon {r} not ready after {timeout}s:\n{self.read_log(r)}")

    def read_log(self, r: int) -> str:
        try:
            return Path(self.logs[r]).read_text(errors="replace")
        except OSError:
            return ""

    def kill(self, rank: int, sig: int = signal.SIGKILL) -> None:
        p = self.procs.get(rank)
        if p and p.poll() is None:
            p.send_signal(sig)
            p.wait(timeout=10)

    def restart(self, rank: int, timeout: float = 60.0) -> None:
        """Kill ``rank``'s daemon (if it still runs) and start a fresh one on the same nodefile line."""
        self.kill(rank)
        log = self.logs[rank] + ".restart"
        self.logs[rank] = log
        with open(log, "w") as f:
            self.procs[rank] = subprocess.Popen([str(_native.DAEMON_PATH), self.nodefile], env=self.env(rank), stdout=f,
                                                stderr=subprocess.STDOUT, start_new_session=True)
        self._wait_ready(rank, timeout)

    def stop(self) -> None:
        for r, p in self.procs.items():
            if p.poll() is None:
                p.send_signal(signal.SIGTERM)
        for r, p in self.procs.items():
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                p.kill()
                p.wait()
        self.procs.clear()
        if self._tmp:
            self._tmp.cleanup()
            self._tmp = None

    def __enter__(self) -> "LocalCluster":
        return self.start()

    def __exit__(self, *exc) -> None:
        self.stop()
