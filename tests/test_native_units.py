"""Native unit suites of the host substrate (tests/cpp/unit_host.cpp), one pytest case per suite.

The reference has no unit layer: wire format, nodefile parser, MPSC ring, lock-free
tables (single- and multi-threaded), governor policies, pmsg mailboxes, socket fd passing, CPU oracles.
"""
import subprocess

import pytest

from conftest import BIN

SUITES = ["msg", "nodefile", "ring", "tables", "tables_mt", "slabs", "frag", "placement", "governor", "topo", "nodefile_numa", "pmsg", "sock", "oracle"]


def test_daemon_check_mode(tmp_path):
    """`oncillamem <nodefile> --check` parses, prints the topology and starts nothing."""
    import os
    import subprocess

    from conftest import BIN, REPO

    env = dict(os.environ, OCM_RANK="5")
    r = subprocess.run([str(BIN / "oncillamem"), str(REPO / "examples" / "nodefile.8gpu"), "--check"], env=env,
                       capture_output=True, text=True, timeout=30)
    assert r.returncode == 0 and "8 rank(s); this process would be rank 5" in r.stdout
    assert r.stdout.count("127.0.0.1") == 8 and "host NUMA nodes:" in r.stdout
    bad = tmp_path / "nodefile"
    bad.write_text("0 a 1.2.3.4 notaport 0\n")
    r = subprocess.run([str(BIN / "oncillamem"), str(bad), "--check"], capture_output=True, text=True, timeout=30)
    assert r.returncode == 1 and "bad ocm_port" in r.stderr


@pytest.mark.parametrize("suite", SUITES)
def test_unit_suite(suite):
    r = subprocess.run([str(BIN / "unit_host"), suite], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert f"{suite}: pass" in r.stdout
