"""Device timing: CUDA events on the launching stream, max over ranks (never wall clock).

(The reference times steps with ``time.time()`` around a loop whose ``.item()`` calls happen to
synchronise the device: pytorch/single_gpu.py:92-94,105-112.  Here timing is explicit.)"""
from __future__ import annotations

import os
import subprocess
import threading
import time

import torch
import torch.distributed as dist


class DeviceTimer:
    def __init__(self, device=None):
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        if self.cuda:
            self.start_ev = torch.cuda.Event(enable_timing=True)
            self.stop_ev = torch.cuda.Event(enable_timing=True)

    def start(self):
        if self.cuda:
            torch.cuda.synchronize()
            self.start_ev.record()
        else:
            self.t0 = time.perf_counter()

    def stop_ms(self) -> float:
        if self.cuda:
            self.stop_ev.record()
            torch.cuda.synchronize()
            return self.start_ev.elapsed_time(self.stop_ev)
        return (time.perf_counter() - self.t0) * 1e3


def max_over_ranks(value: float, group=None) -> float:
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def flush_l2(device=None, nbytes=256 << 20):
    """Writes a buffer larger than the 126 MB L2 so the next kernel starts cold."""
    buf = getattr(flush_l2, "_buf", None)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device or "cuda")
        flush_l2._buf = buf
    buf.zero_()


class ClockSampler:
    """Samples SM clocks and throttle reasons while a timed region runs.

    Uses NVML (the library behind ``nvidia-smi``) from a background thread every ``period_ms``;
    falls back to an ``nvidia-smi --query-gpu=... -lms`` subprocess when NVML cannot be loaded.
    ``summary()`` reports the median SM clock over the samples, the maximum SM clock and the
    set of slowdown reasons that were active at any sample.
    """

    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
    REASON_BITS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
                   0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, gpu_index=0, period_ms=20):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = self._proc = None
        self._nvml = None

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[self.gpu_index])
            except (ValueError, IndexError):
                pass
        return self.gpu_index

    def _nvml_sample(self, handle, nvml):
        self.samples.append(float(nvml.nvmlDeviceGetClockInfo(handle, nvml.NVML_CLOCK_SM)))
        try:
            mask = nvml.nvmlDeviceGetCurrentClocksEventReasons(handle)
        except Exception:  # noqa: BLE001  (older bindings spell it ThrottleReasons)
            mask = nvml.nvmlDeviceGetCurrentClocksThrottleReasons(handle)
        for bit, name in self.REASON_BITS.items():
            if mask & bit:
                self.reasons.add(name)

    def _nvml_loop(self, handle, nvml):
        while not self._stop.is_set():
            try:
                self._nvml_sample(handle, nvml)
            except Exception:  # noqa: BLE001
                break
            self._stop.wait(self.period_ms / 1e3)

    def sample_now(self):
        """One synchronous sample from the calling thread (the timed loop calls this while the GPU is
        busy, so a region too short for the background thread still has a clock record)."""
        if self._nvml is not None:
            try:
                self._nvml_sample(*self._nvml)
            except Exception:  # noqa: BLE001
                pass

    def __enter__(self):
        try:
            import pynvml as nvml
            nvml.nvmlInit()
            handle = nvml.nvmlDeviceGetHandleByIndex(self._physical_index())
            self.max_mhz = float(nvml.nvmlDeviceGetMaxClockInfo(handle, nvml.NVML_CLOCK_SM))
            self._nvml = (handle, nvml)
            self._thread = threading.Thread(target=self._nvml_loop, args=(handle, nvml), daemon=True)
            self._thread.start()
            return self
        except Exception:  # noqa: BLE001
            pass
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-i", str(self._physical_index()), "-lms", str(max(self.period_ms, 50))],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._thread = threading.Thread(target=self._smi_loop, daemon=True)
            self._thread.start()
        except OSError:
            self._proc = None
        return self

    def _smi_loop(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self._proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                self.samples.append(float(parts[0]))
                self.max_mhz = max(self.max_mhz or 0.0, float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    self.reasons.add(name)

    def __exit__(self, *exc):
        self._stop.set()
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)

    def summary(self):
        sm = sorted(self.samples)
        med = sm[len(sm) // 2] if sm else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(sm)}
