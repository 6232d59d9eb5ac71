"""Native unit suites of the host substrate (tests/cpp/unit_host.cpp), one pytest case per suite.

The reference has no unit layer: wire format, nodefile parser, MPSC ring, lock-free
tables (single- and multi-threaded), governor policies, pmsg mailboxes, socket fd passing, CPU oracles.
"""
import subprocess

import pytest

from conftest import BIN

SUITES = ["msg", "nodefile", "ring", "tables", "tables_mt", "governor", "pmsg", "sock", "oracle"]


@pytest.mark.parametrize("suite", SUITES)
def test_unit_suite(suite):
    r = subprocess.run([str(BIN / "unit_host"), suite], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert f"{suite}: pass" in r.stdout
