"""Samples `nvidia-smi` clocks and throttle reasons while a timed region runs
(B200_PROFILING.md: a number taken under hw/thermal slowdown is rejected)."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")
_REASONS = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")


class ClockSampler:
    """with ClockSampler(gpu_index) as cs: <timed region>; cs.summary() -> dict"""

    def __init__(self, gpu_index: int = 0, period_ms: int = 50):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.samples: List[dict] = []
        self._proc: Optional[subprocess.Popen] = None
        self._thread: Optional[threading.Thread] = None

    def _reader(self) -> None:
        assert self._proc is not None and self._proc.stdout is not None
        for line in self._proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 8 or not parts[0].isdigit():
                continue
            try:
                self.samples.append({
                    "t": time.monotonic(), "sm": float(parts[1].split()[0]),
                    "max": float(parts[2].split()[0]),
                    "power": float(parts[3].split()[0]) if parts[3][0].isdigit() else 0.0,
                    "reasons": [n for n, v in zip(_REASONS, parts[4:8]) if v.lower() == "active"],
                })
            except (ValueError, IndexError):
                continue

    def __enter__(self) -> "ClockSampler":
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader",
                 "-i", str(self.gpu_index), "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._thread = threading.Thread(target=self._reader, daemon=True)
            self._thread.start()
            # nvidia-smi needs a moment to come up; wait for its first line so that even a
            # timed region of a few tens of ms is bracketed by samples
            deadline = time.monotonic() + 3.0
            while not self.samples and time.monotonic() < deadline:
                time.sleep(0.01)
        except OSError:
            self._proc = None
        self.t0 = time.monotonic()
        return self

    def __exit__(self, *exc) -> None:
        self.t1 = time.monotonic()
        if self._proc is not None:
            time.sleep(self.period_ms / 1000.0 * 1.5)  # let the last sample land
            self._proc.terminate()
            try:
                self._proc.wait(2)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            if self._thread is not None:
                self._thread.join(2)

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        inside = [s for s in self.samples if self.t0 - 0.06 <= s["t"] <= self.t1 + 0.3] or self.samples
        reasons = sorted({r for s in inside for r in s["reasons"]})
        return {"sm_mhz": statistics.median(s["sm"] for s in inside),
                "sm_max_mhz": max(s["max"] for s in inside),
                "power_w_max": max(s["power"] for s in inside),
                "reasons": reasons, "samples": len(inside)}
