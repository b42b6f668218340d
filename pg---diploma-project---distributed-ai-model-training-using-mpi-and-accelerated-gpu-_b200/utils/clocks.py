"""GPU clock / throttle sampling during a timed region (B200_PROFILING.md "clocks line").

``ClockSampler`` runs ``nvidia-smi -lms 200`` in the background between ``start()`` and ``stop()``
and summarises median SM clock under load, max clock and active throttle reasons, so a bench line
can show the run was not thermally or externally throttled.  Clocks are never *set* here.
"""
from __future__ import annotations

import shutil
import statistics
import subprocess
import tempfile
from typing import Any, Dict, List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
_REASONS = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")


class ClockSampler:
    def __init__(self, gpu_index: Optional[int] = None, period_ms: int = 200) -> None:
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._file = None

    def start(self) -> "ClockSampler":
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self._file = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        cmd = [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-lms", str(self.period_ms)]
        if self.gpu_index is not None:
            cmd += ["-i", str(self.gpu_index)]
        try:
            self._proc = subprocess.Popen(cmd, stdout=self._file, stderr=subprocess.DEVNULL)
        except OSError:
            self._proc = None
        return self

    def stop(self) -> Dict[str, Any]:
        if self._proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0,
                    "note": "nvidia-smi unavailable"}
        self._proc.terminate()      # exact PID we started
        try:
            self._proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self._proc.kill()
        self._file.flush()
        self._file.seek(0)
        return summarize_clock_csv(self._file.read().splitlines())


def summarize_clock_csv(lines: List[str]) -> Dict[str, Any]:
    sm: List[float] = []
    mx: List[float] = []
    pw: List[float] = []
    reasons = set()
    for ln in lines:
        parts = [p.strip() for p in ln.split(",")]
        if len(parts) < 9:
            continue
        try:
            sm.append(float(parts[1]))
            mx.append(float(parts[2]))
        except ValueError:
            continue
        try:
            pw.append(float(parts[3]))
        except ValueError:
            pass
        for name, val in zip(_REASONS, parts[5:9]):
            if val.lower().startswith("active"):
                reasons.add(name)
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
    # "under load" = samples drawing more than the idle floor; fall back to all samples
    load = [c for c, p in zip(sm, pw)] if len(pw) != len(sm) else [c for c, p in zip(sm, pw) if p > 250.0]
    use = load if load else sm
    return {"sm_mhz": statistics.median(use), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
            "samples": len(sm), "power_w_max": max(pw) if pw else None}
