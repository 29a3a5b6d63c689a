"""Measurement helpers: CUDA-event timers (max over ranks), an L2 flush, and an nvidia-smi clock
sampler that runs DURING the timed region (B200_PROFILING.md "clocks line")."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import Dict, List, Optional

import torch

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples SM clocks / throttle reasons every ``period_ms`` in a background process."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._lines: List[str] = []
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "ClockSampler":
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-lms", str(self.period_ms),
                 "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self

        def pump():
            assert self._proc is not None and self._proc.stdout is not None
            for line in self._proc.stdout:
                self._lines.append(line.strip())

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict[str, object]:
        if self._proc is not None:
            time.sleep(self.period_ms / 1000.0 * 1.5)
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            if self._thread is not None:
                self._thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for line in self._lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                mx.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 between timed iterations."""

    def __init__(self, device: torch.device, mbytes: int = 256):
        self.buf = torch.empty(mbytes << 20, dtype=torch.uint8, device=device)

    def flush(self) -> None:
        self.buf.zero_()


def max_over_ranks(value: float, world: int) -> float:
    """Max of a host float over torch.distributed ranks (any backend); identity when world == 1."""
    if world <= 1:
        return value
    import torch.distributed as dist
    if not dist.is_initialized():
        return value
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
