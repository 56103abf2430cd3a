"""Sample SM clocks and throttle reasons while a timed region runs (B200 profiling recipe)."""

from __future__ import annotations

import shutil
import statistics
import subprocess
import tempfile

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
_REASONS = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


class ClockSampler:
    """``with ClockSampler(gpu_index) as c: ...`` then ``c.summary()``."""

    def __init__(self, index: int = 0, period_ms: int = 100):
        self.index, self.period_ms = index, period_ms
        self._proc = None
        self._file = None

    def __enter__(self):
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self._file = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self._proc = subprocess.Popen(
            [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.index),
             "-lms", str(self.period_ms)], stdout=self._file, stderr=subprocess.DEVNULL)
        return self

    def __exit__(self, *exc):
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:  # pragma: no cover
                self._proc.kill()
        return False

    def summary(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self._file is None:
            out["reasons"] = ["nvidia-smi unavailable"]
            return out
        self._file.flush()
        self._file.seek(0)
        sm, mx, power = [], [], []
        reasons = set()
        for line in self._file.read().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(_REASONS, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            # "under load": samples in the upper half of the power range
            thr = (max(power) + min(power)) / 2 if power else 0
            loaded = [c for c, pw in zip(sm, power) if pw >= thr] or sm
            out.update(sm_mhz=statistics.median(loaded), sm_max_mhz=max(mx), samples=len(sm),
                       power_w_max=max(power))
        out["reasons"] = sorted(reasons)
        return out
