"""Device-side timing and clock sampling (the reference has one wall clock)."""
import statistics
import subprocess
import threading

import torch

_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples nvidia-smi every `period_ms` in the background while a timed region runs."""

    def __init__(self, gpu_index=0, period_ms=200):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.samples, self._proc, self._thread = [], None, None

    def __enter__(self):
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index),
                 "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self

        def pump():
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) >= 7:
                    self.samples.append(parts)
        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            self._thread.join(timeout=5)
        return False

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx.append(float(s[1]))
            except ValueError:
                continue
            for name, val in zip(names, s[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def time_steps(step_fn, steps, warmup, barrier=None):
    """Run `warmup` untimed then exactly `steps` timed calls of step_fn on the current
    stream; CUDA events around the timed region with a synchronize on both sides."""
    for _ in range(warmup):
        step_fn()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    torch.cuda.synchronize()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(steps):
        step_fn()
    stop.record()
    torch.cuda.synchronize()
    if barrier:
        barrier()
    return start.elapsed_time(stop) / 1e3  # seconds
