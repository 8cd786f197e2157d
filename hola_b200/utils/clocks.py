"""Machine-state sampling during a timed region.

The bench contract asks for GPU clocks and throttle reasons sampled with nvidia-smi while the
measurement runs. This workload never touches the GPU, so the GPU line is
recorded only as box context; what can actually move a CPU-bound HTTP number is the CPU clock,
which is sampled from /proc/cpuinfo alongside.
"""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional

_QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
_REASONS = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]


def cpu_info() -> Dict:
    model, mhz = None, []
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name") and model is None:
                model = line.split(":", 1)[1].strip()
            elif line.startswith("cpu MHz"):
                try:
                    mhz.append(float(line.split(":", 1)[1]))
                except ValueError:
                    pass
    except OSError:
        pass
    return {"model": model, "mhz_median": round(statistics.median(mhz), 1) if mhz else None, "logical_cpus": len(mhz) or None}


class ClockSampler:
    """Samples nvidia-smi (if present) and /proc/cpuinfo every ``period`` seconds in a thread."""

    def __init__(self, period: float = 2.0):
        self.period = period
        self.smi = shutil.which("nvidia-smi")
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self.sm_mhz: List[float] = []
        self.sm_max_mhz: Optional[float] = None
        self.reasons: set = set()
        self.cpu_mhz: List[float] = []

    def _sample_gpu(self) -> None:
        if not self.smi:
            return
        try:
            # nvidia-smi is a heavy process (tens of ms of CPU per call) and every CPU of the box is
            # busy with the pinned server and generator threads: run it at the lowest priority so
            # that it fills gaps instead of stalling a loop that is being measured.
            out = subprocess.run([self.smi, "--query-gpu=" + _QUERY, "--format=csv,noheader,nounits", "-i", "0"],
                                 stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=10,
                                 preexec_fn=lambda: os.nice(19)).stdout
        except (OSError, subprocess.TimeoutExpired):
            return
        for line in out.strip().splitlines():
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 2 + len(_REASONS):
                continue
            try:
                self.sm_mhz.append(float(parts[0]))
                self.sm_max_mhz = float(parts[1])
            except ValueError:
                continue
            for name, value in zip(_REASONS, parts[2:]):
                if value.lower().startswith("active"):
                    self.reasons.add(name)

    def _run(self) -> None:
        while not self._stop.is_set():
            self._sample_gpu()
            mhz = cpu_info()["mhz_median"]
            if mhz:
                self.cpu_mhz.append(mhz)
            self._stop.wait(self.period)

    def start(self) -> "ClockSampler":
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict:
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=10)
        gpu = None
        if self.sm_mhz:
            gpu = {"sm_mhz": statistics.median(self.sm_mhz), "sm_max_mhz": self.sm_max_mhz, "reasons": sorted(self.reasons)}
        return {
            "sm_mhz": gpu["sm_mhz"] if gpu else None,
            "sm_max_mhz": gpu["sm_max_mhz"] if gpu else None,
            "reasons": gpu["reasons"] if gpu else [],
            "gpu_sampled": bool(gpu),
            "note": "GPU idle: this workload is CPU-only; GPU clocks are box context, not a factor in the number",
            "cpu_mhz_median": round(statistics.median(self.cpu_mhz), 1) if self.cpu_mhz else None,
        }
