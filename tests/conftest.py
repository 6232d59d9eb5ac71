"""pytest configuration.

Markers: ``gpu`` - needs a real B200 (run by ``pytest -m gpu`` on the GPU box).  Everything else runs on
the CPU-only build machine: native unit tests, daemon/protocol integration with several daemons on one
host, allocator stress, oracles.
"""
import os
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
BIN = REPO / "build" / "bin"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA GPU (B200)")
    if not (REPO / "build" / "lib" / "libocm.so").exists():
        import __graft_entry__

        __graft_entry__.build()


def have_gpu() -> bool:
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if have_gpu():
        return
    skip = pytest.mark.skip(reason="no GPU on this machine")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture
def cluster2():
    """Two CPU-only ranks on this host."""
    import oncilla_b200 as ocm

    with ocm.LocalCluster(2, host_bytes=32 << 20, extra_env={"OCM_NO_GPU": "1"}) as cl:
        yield cl


def run_app(cl, rank, argv, timeout=120, env_extra=None):
    """Run a native test program attached to ``rank``'s daemon."""
    env = cl.env(rank)
    if env_extra:
        env.update(env_extra)
    return subprocess.run([str(a) for a in argv], env=env, capture_output=True, text=True, timeout=timeout)


def run_py(cl, rank, code, timeout=180, env_extra=None, argv=()):
    """Run a Python snippet in a fresh process attached to ``rank``'s daemon."""
    env = cl.env(rank)
    env["PYTHONPATH"] = str(REPO) + os.pathsep + env.get("PYTHONPATH", "")
    if env_extra:
        env.update(env_extra)
    return subprocess.run([sys.executable, "-c", code, *argv], env=env, capture_output=True, text=True, timeout=timeout)


def pytest_terminal_summary(terminalreporter, exitstatus, config):
    """One line that names every skipped test: a GPU lease with fewer devices than a test needs must show up in the
    log as a visible gap, not as a silent 's'."""
    skipped = terminalreporter.stats.get("skipped", [])
    if not skipped:
        return
    names = sorted({r.nodeid.split("::", 1)[-1] for r in skipped})
    why = sorted({str(r.longrepr[-1]) if isinstance(r.longrepr, tuple) else str(r.longrepr) for r in skipped})
    terminalreporter.write_line(f"SKIPPED ({len(names)}): " + ", ".join(names))
    terminalreporter.write_line("SKIP REASONS: " + " | ".join(w[:120] for w in why))
