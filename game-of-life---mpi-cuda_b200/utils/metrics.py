"""Throughput and roofline arithmetic."""
from __future__ import annotations

import json
import os
from typing import Optional


def gcups(cell_updates: float, seconds: float) -> float:
    """Giga cell updates per second = nranks*H*W*iterations / duration (gol-main.c:124-125)."""
    return cell_updates / seconds * 1e-9 if seconds > 0 else 0.0


def measured_peaks(repo_root: Optional[str] = None) -> dict:
    root = repo_root or os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    path = os.path.join(root, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        d["source"] = "measured"
        return d
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback"}


def roofline_fraction(cells_per_s_per_gpu: float, generations_per_pass: int = 1,
                      hbm_gbs: Optional[float] = None) -> dict:
    """Fraction of the bit-packed HBM roofline: 1 bit read + 1 bit written per cell per pass."""
    peaks = measured_peaks()
    bw = (hbm_gbs or peaks["hbm_gbs"]) * 1e9
    bytes_per_cell_gen = 0.25 / max(1, generations_per_pass)
    bound = bw / bytes_per_cell_gen
    return {"bound_cells_per_s": bound, "fraction": cells_per_s_per_gpu / bound, "of": peaks["source"]}
