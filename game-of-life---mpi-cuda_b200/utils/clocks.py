"""Sample SM clocks and throttle reasons with nvidia-smi while a timed region runs."""
from __future__ import annotations

import shutil
import statistics
import subprocess
import threading
import time
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
                    "t": time.time(),
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

    def wait_first_sample(self, timeout_s: float = 10.0) -> bool:
        """nvidia-smi needs a second or more before its first line arrives (longer with eight ranks
        starting at once): block until it has, so that a short timed region is not left without a
        clock record."""
        t_end = time.time() + timeout_s
        while not self.samples and time.time() < t_end:
            if self._proc is None or self._proc.poll() is not None:
                return False
            time.sleep(0.05)
        return bool(self.samples)

    def stop(self, t0: Optional[float] = None, t1: Optional[float] = None) -> Dict:
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            if self._thread is not None:
                self._thread.join(timeout=2)
        return self.summary(t0, t1)

    def summary(self, t0: Optional[float] = None, t1: Optional[float] = None) -> Dict:
        """Median SM clock, max clock and throttle reasons.  With a window [t0, t1] (time.time())
        only the samples taken inside it count; if the window caught none (very short region) the
        samples under load (highest power quartile) of the whole run are used and flagged."""
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        chosen, scope = self.samples, "whole run"
        if t0 is not None and t1 is not None:
            inside = [s for s in self.samples if t0 <= s["t"] <= t1]
            if inside:
                chosen, scope = inside, "timed region"
            else:
                cut = sorted(s["power"] for s in self.samples)[(3 * len(self.samples)) // 4]
                chosen, scope = [s for s in self.samples if s["power"] >= cut], "under load (region too short)"
        sm = [s["sm"] for s in chosen]
        reasons = sorted({r for s in chosen for r in s["reasons"]})
        return {
            "sm_mhz": statistics.median(sm),
            "sm_max_mhz": max(s["max"] for s in chosen),
            "power_w_max": max(s["power"] for s in chosen),
            "reasons": reasons,
            "samples": len(chosen),
            "scope": scope,
        }

    def __enter__(self):
        return self.start()

    def __exit__(self, *exc):
        self.stop()
