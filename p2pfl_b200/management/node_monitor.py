"""Periodic resource sampler.

Host counters as in the reference (cpu %, ram %, network MiB/s;
p2pfl/management/node_monitor.py:58-82) plus per-GPU counters through NVML
when it is available (memory used, SM clock, power, utilisation), because on
an 8xB200 box those are the interesting ones.
"""

from __future__ import annotations

import threading
import time
from typing import Callable, Dict, Optional

from p2pfl_b200.settings import Settings

try:  # psutil is present in the image but keep the monitor optional
    import psutil
except Exception:  # pragma: no cover
    psutil = None  # type: ignore

ReportFn = Callable[[str, str, float], None]


class NodeMonitor(threading.Thread):
    """Daemon thread calling ``report(node, metric, value)`` every period."""

    def __init__(self, node_addr: str, report: ReportFn, gpu_index: Optional[int] = None) -> None:
        super().__init__(name=f"resource-monitor-thread-{node_addr}", daemon=True)
        self.node_addr = node_addr
        self.report = report
        self.gpu_index = gpu_index
        self._stop_evt = threading.Event()
        self._nvml = None
        self._last_net: Optional[tuple] = None

    def stop(self) -> None:
        self._stop_evt.set()

    # -- samplers ----------------------------------------------------------
    def sample_host(self) -> Dict[str, float]:
        if psutil is None:
            return {}
        out = {"cpu": float(psutil.cpu_percent()), "ram": float(psutil.virtual_memory().percent)}
        net = psutil.net_io_counters()
        now = time.time()
        if self._last_net is not None:
            t0, rx0, tx0 = self._last_net
            dt = max(now - t0, 1e-6)
            out["net_in"] = (net.bytes_recv - rx0) / dt / (1 << 20)
            out["net_out"] = (net.bytes_sent - tx0) / dt / (1 << 20)
        self._last_net = (now, net.bytes_recv, net.bytes_sent)
        return out

    def sample_gpu(self) -> Dict[str, float]:
        if self.gpu_index is None:
            return {}
        try:
            import pynvml

            if self._nvml is None:
                pynvml.nvmlInit()
                self._nvml = pynvml.nvmlDeviceGetHandleByIndex(self.gpu_index)
            h = self._nvml
            mem = pynvml.nvmlDeviceGetMemoryInfo(h)
            util = pynvml.nvmlDeviceGetUtilizationRates(h)
            return {
                "gpu_mem_gb": mem.used / 1e9,
                "gpu_util": float(util.gpu),
                "gpu_sm_mhz": float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)),
                "gpu_power_w": pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0,
            }
        except Exception:
            self.gpu_index = None  # NVML not usable here; stop trying
            return {}

    def run(self) -> None:
        while not self._stop_evt.wait(Settings.RESOURCE_MONITOR_PERIOD):
            for k, v in {**self.sample_host(), **self.sample_gpu()}.items():
                self.report(self.node_addr, k, v)
