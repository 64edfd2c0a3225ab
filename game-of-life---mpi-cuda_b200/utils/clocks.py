"""Sample SM clocks and throttle reasons with nvidia-smi while a timed region runs."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
from typing import Dict, List, Optional

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
_REASONS = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


class ClockSampler:
    def __init__(self, gpu_index: Optional[int] = None, period_ms: int = 100):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.samples: List[Dict] = []
        self._proc = None
        self._thread = None

    def _reader(self):
        for line in self._proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9 or not parts[0].isdigit():
                continue
            try:
                self.samples.append({
                    "index": int(parts[0]),
                    "sm": float(parts[1].split()[0]),
                    "max": float(parts[2].split()[0]),
                    "power": float(parts[3].split()[0]) if parts[3][0].isdigit() else 0.0,
                    "reasons": [r for r, v in zip(_REASONS, parts[5:9]) if v.lower().startswith("active")],
                })
            except (ValueError, IndexError):
                continue

    def start(self) -> "ClockSampler":
        exe = shutil.which("nvidia-smi")
        if not exe:
            return self
        cmd = [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader", "-lms", str(self.period_ms)]
        if self.gpu_index is not None:
            cmd += ["-i", str(self.gpu_index)]
        self._proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        self._thread = threading.Thread(target=self._reader, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict:
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            if self._thread is not None:
                self._thread.join(timeout=2)
        return self.summary()

    def summary(self) -> Dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [s["sm"] for s in self.samples]
        reasons = sorted({r for s in self.samples for r in s["reasons"]})
        return {
            "sm_mhz": statistics.median(sm),
            "sm_max_mhz": max(s["max"] for s in self.samples),
            "power_w_max": max(s["power"] for s in self.samples),
            "reasons": reasons,
            "samples": len(self.samples),
        }

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()
