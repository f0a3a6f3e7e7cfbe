"""Device-side timing and clock sampling (the reference has one wall clock)."""
import os
import statistics
import subprocess
import threading

import torch

_QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
          "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
          "clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Samples SM clocks and throttle reasons of one GPU in a background thread while a
    timed region runs.  Uses NVML in-process (first sample within milliseconds, so short
    regions are covered); falls back to polling nvidia-smi."""

    _REASONS = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, gpu_index=0, period_ms=50):
        self.gpu_index, self.period = gpu_index, period_ms / 1e3
        self.sm, self.reasons, self.sm_max = [], set(), None
        self._stop = threading.Event()
        self._thread = None

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if self.gpu_index < len(ids) and ids[self.gpu_index].isdigit():
                return int(ids[self.gpu_index])
        return self.gpu_index

    def _run_nvml(self):
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index())
        self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
            pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop.is_set():
            self.sm.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
            mask = int(get_reasons(h))
            for name, bit in self._REASONS.items():
                if mask & bit:
                    self.reasons.add(name)
            self._stop.wait(self.period)

    def _run_smi(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-i",
                                      str(self._physical_index())], capture_output=True, text=True, timeout=10).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                self.sm.append(float(parts[0]))
                self.sm_max = float(parts[1])
                for name, val in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"],
                                     parts[3:7]):
                    if val.lower().startswith("active"):
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(max(self.period, 0.2))

    def _run(self):
        try:
            self._run_nvml()
        except Exception:
            self._run_smi()

    def __enter__(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thread.join(timeout=15)
        return False

    def summary(self):
        return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.sm_max,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


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


class PhaseTracer:
    """Per-phase device timing + NVTX ranges (the reference has one wall clock).

    with tracer.phase("pack"): ...   NVTX range always; CUDA events only when enabled.
    report() -> {phase: milliseconds} of the last step (synchronises)."""

    def __init__(self, enabled=False):
        self.enabled = enabled
        self._events = []

    class _Phase:
        def __init__(self, tracer, name):
            self.t, self.name = tracer, name

        def __enter__(self):
            torch.cuda.nvtx.range_push(self.name)
            if self.t.enabled:
                self.e0 = torch.cuda.Event(enable_timing=True)
                self.e0.record()
            return self

        def __exit__(self, *exc):
            if self.t.enabled:
                e1 = torch.cuda.Event(enable_timing=True)
                e1.record()
                self.t._events.append((self.name, self.e0, e1))
            torch.cuda.nvtx.range_pop()
            return False

    def phase(self, name):
        return PhaseTracer._Phase(self, name)

    def begin_step(self):
        self._events = []

    def report(self):
        if not self.enabled:
            return {}
        torch.cuda.synchronize()
        out = {}
        for name, e0, e1 in self._events:
            out[name] = out.get(name, 0.0) + e0.elapsed_time(e1)
        return out
