"""Measurement helpers: clock sampling during timed regions, device timing, L2 flush."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
import time

_Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
      "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
      "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples ``nvidia-smi`` every 200 ms on a background thread (B200_PROFILING.md clocks line)."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu = gpu_index
        self.period_ms = period_ms
        self.samples: list[list[str]] = []
        self._proc = None
        self._thr = None

    def __enter__(self):
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self._proc = subprocess.Popen([exe, f"--query-gpu={_Q}", "--format=csv,noheader,nounits", "-i", str(self.gpu),
                                       "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                                      text=True)

        def pump():
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) >= 9:
                    self.samples.append(parts)

        self._thr = threading.Thread(target=pump, daemon=True)
        self._thr.start()
        return self

    def __exit__(self, *exc):
        if self._proc:
            time.sleep(0.05)
            self._proc.terminate()
            try:
                self._proc.wait(timeout=3)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            if self._thr:
                self._thr.join(timeout=2)

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm, mx = [], []
        reasons = set()
        for p in self.samples:
            try:
                sm.append(float(p[1]))
                mx.append(float(p[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 between timed iterations."""

    def __init__(self, nbytes: int = 256 << 20):
        import torch

        self.buf = torch.empty(nbytes, dtype=torch.uint8, device="cuda")

    def flush(self):
        self.buf.zero_()
