"""The driver-facing contract of bench.py and __graft_entry__.py, as far as a machine without a GPU can check it."""
import json
import subprocess
import sys

import pytest

from conftest import REPO


def test_reference_arm_prints_one_json_line_and_exits_zero():
    """`bench.py --impl reference` must never fail the driver: without a GPU (or without the built reference)
    it reports why the arm is unavailable."""
    r = subprocess.run([sys.executable, str(REPO / "bench.py"), "--impl", "reference", "--steps", "2", "--warmup", "3"],
                       capture_output=True, text=True, timeout=300, cwd=str(REPO))
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    out = json.loads(lines[0])
    assert out["impl"] == "reference"
    assert "unavailable" in out or out.get("metric", "").startswith("ocm_copy GB/s")


def test_both_arms_print_the_same_metric_string():
    """The driver only computes vs_reference when metric, unit and direction agree: one constant serves both arms."""
    sys.path.insert(0, str(REPO / "baseline"))
    import ref_harness

    src = (REPO / "bench.py").read_text()
    assert "from ref_harness import METRIC" in src and '"metric": METRIC' in src
    assert '"metric": METRIC' in (REPO / "baseline" / "ref_harness.py").read_text()
    assert ref_harness.METRIC.startswith("ocm_copy GB/s")


def test_bench_defaults_and_flags():
    r = subprocess.run([sys.executable, str(REPO / "bench.py"), "--help"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in r.stdout


def test_graft_entry_has_build_and_smoke():
    sys.path.insert(0, str(REPO))
    import __graft_entry__ as g

    assert callable(g.build) and callable(g.smoke)


def test_checkpoint_header_rejects_garbage(tmp_path):
    """Files that are not checkpoints (or are cut off / tampered with in the header) are refused before any data moves."""
    import struct

    from oncilla_b200 import checkpoint as ckpt

    cases = {
        "empty": b"",
        "wrong_magic": b"NOTOCMCK" + struct.pack("<Q", 10) + b"{}" * 5,
        "huge_header": ckpt.MAGIC + struct.pack("<Q", 1 << 40),
        "not_json": ckpt.MAGIC + struct.pack("<Q", 6) + b"\xff\xfe{{{{",
        "missing_keys": ckpt.MAGIC + struct.pack("<Q", 2) + b"{}",
    }
    for name, blob in cases.items():
        p = tmp_path / name
        p.write_bytes(blob)
        with pytest.raises(ckpt.CheckpointError):
            ckpt.read_header(str(p))


@pytest.mark.gpu
def test_bench_json_contract_on_gpu():
    """One short device-timed run: every key the driver reads is present and sane."""
    r = subprocess.run([sys.executable, str(REPO / "bench.py"), "--steps", "3", "--warmup", "3", "--curve-max", str(16 << 20)],
                       capture_output=True,
                       text=True, timeout=600, cwd=str(REPO))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    out = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches"):
        assert key in out, key
    assert out["n_gpus"] == 1 and out["steps"] == 3 and out["warmup"] >= 3 and out["higher_is_better"] is True
    assert out["value"] > 100 and out["gpu_launches"] > 0
    assert out["e2e"]["value"] > 1 and out["e2e"]["h2d_bytes_per_step"] > 0 and out["e2e"]["d2h_bytes_per_step"] > 0
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(out["clocks"])
    sys.path.insert(0, str(REPO / "baseline"))
    import ref_harness

    assert out["metric"] == ref_harness.METRIC and out["unit"] == "GB/s"
    assert [r["bytes"] for r in out["curve"]] == [4 << 10, 64 << 10, 1 << 20, 16 << 20]
    for r in out["curve"]:
        assert r["put_GBps"] > 0 and r["get_GBps"] > 0 and r["copy_GBps"] > 0
