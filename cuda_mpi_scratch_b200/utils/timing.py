"""Device-side timing and clock sampling.

The reference times with ``clock()`` (CPU time, compared across processes,
``mpicuda3.cu:176-179,315-325``) and a single un-warmed ``MPI_Wtime`` bracket
(``test-benchmark/mpi-pingpong-gpu.cpp:51-54``) -- SURVEY.md 2.7 #12.  Here
every number is taken with CUDA events on the launching stream after warm-up,
and multi-rank numbers are the max over ranks.
"""
from __future__ import annotations

import statistics
import subprocess
import threading
from typing import Callable

import torch

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 between timed iterations."""

    def __init__(self, device: int, nbytes: int = 256 << 20):
        self.buf = torch.empty(nbytes // 4, dtype=torch.int32, device=torch.device("cuda", device))
        self._v = 0

    def __call__(self) -> None:
        self._v += 1
        self.buf.fill_(self._v)


def time_cuda_ms(fn: Callable[[], None], warmup: int = 3, iters: int = 10, device: int = 0,
                 flush: L2Flusher | None = None) -> list[float]:
    """Per-iteration device time of ``fn`` in ms (CUDA events, after warm-up)."""
    stream = torch.cuda.current_stream(device)
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(device)
    out = []
    for _ in range(iters):
        if flush is not None:
            flush()
        a = torch.cuda.Event(enable_timing=True)
        b = torch.cuda.Event(enable_timing=True)
        a.record(stream)
        fn()
        b.record(stream)
        b.synchronize()
        out.append(a.elapsed_time(b))
    return out


class ClockSampler:
    """Samples ``nvidia-smi`` clocks/throttle reasons in the background while a
    timed region runs (the recipe in B200_PROFILING.md)."""

    def __init__(self, device: int = 0, period_ms: int = 100):
        self.device, self.period_ms = device, period_ms
        self.rows: list[list[str]] = []
        self._proc: subprocess.Popen | None = None
        self._thr: threading.Thread | None = None

    # NVML bit masks of nvmlDeviceGetCurrentClocksEventReasons
    _NVML_REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
                     0x40: "hw_thermal_slowdown"}

    def _start_nvml(self) -> bool:
        """Poll NVML in-process every ~2 ms: dozens of samples even for a 20 ms timed region."""
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self._nvml_index())
            smax = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        except Exception:
            return False
        self._stop_evt = threading.Event()

        def pump():
            import time
            while not self._stop_evt.is_set():
                try:
                    sm = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                    pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                    try:
                        mask = pynvml.nvmlDeviceGetCurrentClocksEventReasons(h)
                    except Exception:
                        mask = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    flags = ["Active" if mask & b else "Not Active" for b in (0x8, 0x40, 0x20, 0x4)]
                    self.rows.append([str(self.device), str(sm), str(smax), f"{pw:.2f}", hex(mask)] + flags)
                except Exception:
                    pass
                time.sleep(0.002)

        self._thr = threading.Thread(target=pump, daemon=True)
        self._thr.start()
        self.source = "nvml"
        return True

    def _nvml_index(self) -> int:
        """NVML enumerates physical GPUs; honour CUDA_VISIBLE_DEVICES when it lists indices."""
        import os
        vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        ids = [v for v in vis.split(",") if v.strip().isdigit()]
        return int(ids[self.device]) if self.device < len(ids) else self.device

    def start(self) -> "ClockSampler":
        self.source = "nvidia-smi"
        if self._start_nvml():
            return self
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits",
                 "-i", str(self.device), "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self._proc = None
            return self
        self._thr = threading.Thread(target=self._pump, daemon=True)
        self._thr.start()
        return self

    def wait_first(self, timeout_s: float = 2.0) -> None:
        """Block until the first sample arrived (nvidia-smi takes a while to start)."""
        import time
        t0 = time.perf_counter()
        while not self.rows and time.perf_counter() - t0 < timeout_s:
            time.sleep(0.01)

    def _pump(self) -> None:
        assert self._proc is not None and self._proc.stdout is not None
        for line in self._proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 9:
                self.rows.append(parts)

    def stop(self) -> dict:
        if getattr(self, "_stop_evt", None) is not None:
            self._stop_evt.set()
        if self._proc is not None:
            self._proc.terminate()
            try:
                self._proc.wait(timeout=2)
            except subprocess.TimeoutExpired:  # pragma: no cover
                self._proc.kill()
        if self._thr is not None:
            self._thr.join(timeout=2)
        return self.summary()

    def window_begin(self) -> None:
        """Mark the start of a timed region: only samples inside windows are summarised."""
        self._windows = getattr(self, "_windows", [])
        self._windows.append([len(self.rows), None])

    def window_end(self) -> None:
        if getattr(self, "_windows", None) and self._windows[-1][1] is None:
            self._windows[-1][1] = len(self.rows)

    def pop_window_summary(self) -> dict:
        """Summary of the LAST closed window only, which is then forgotten: one clock record per
        measurement when a script times many things with one sampler (``bench/*.py``)."""
        wins = getattr(self, "_windows", [])
        if not wins or wins[-1][1] is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        a, b = wins.pop()
        # a window shorter than one sampling period still gets the sample that follows it
        rows = self.rows[a:max(b, a + 1)] or self.rows[-1:]
        return self._summarise(rows, True)

    def summary(self) -> dict:
        rows = self.rows
        wins = [w for w in getattr(self, "_windows", []) if w[1] is not None and w[1] > w[0]]
        in_window = bool(wins)
        if in_window:
            rows = [r for a, b in wins for r in self.rows[a:b]]
        return self._summarise(rows, in_window)

    def _summarise(self, rows, in_window: bool) -> dict:
        sm, smax, power = [], [], []
        reasons: set[str] = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[1]))
                smax.append(float(r[2]))
                power.append(float(r[3]))
            except ValueError:
                continue
            for name, val in zip(names, r[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if in_window:
            load = sm      # every sample was taken inside a timed region
        else:              # "under load": samples in the upper half of the power range
            pmid = (max(power) + min(power)) / 2
            load = [s for s, p in zip(sm, power) if p >= pmid] or sm
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(smax),
                "reasons": sorted(reasons), "samples": len(sm), "samples_in_timed_regions": in_window,
                "power_w_max": max(power), "source": getattr(self, "source", "nvidia-smi")}
