"""Phase timers, Chrome-trace timeline, clock sampling and L2 flushing.

``trainer.get_training_time()`` is the reference's only timer (driver wall
clock, /root/reference/ddl_mnist_aztk.py:222).  Here every throughput number is
timed on the device with CUDA events; a Horovod-timeline-style Chrome trace
can be written for host-side phases; ``ClockSampler`` records SM clocks and
throttle reasons with ``nvidia-smi`` while a timed region runs."""
from __future__ import annotations

import json
import os
import subprocess
import threading
import time
from contextlib import contextmanager
from typing import Dict, List, Optional

import torch


class PhaseTimers:
    """Accumulating host-side phase timers (feed / fwd / bwd / allreduce / opt)."""

    def __init__(self):
        self.total: Dict[str, float] = {}
        self.count: Dict[str, int] = {}

    @contextmanager
    def phase(self, name: str):
        t0 = time.perf_counter()
        try:
            yield
        finally:
            dt = time.perf_counter() - t0
            self.total[name] = self.total.get(name, 0.0) + dt
            self.count[name] = self.count.get(name, 0) + 1

    def summary(self) -> Dict[str, Dict[str, float]]:
        return {k: {'total_s': v, 'count': self.count[k], 'mean_ms': 1e3 * v / self.count[k]}
                for k, v in self.total.items()}


class Timeline:
    """Chrome-trace (``chrome://tracing``) writer, enabled by ``DDLS_TIMELINE=<path>``."""

    def __init__(self, path: Optional[str] = None, rank: int = 0):
        self.path = path or os.environ.get('DDLS_TIMELINE')
        self.rank = rank
        self.events: List[dict] = []
        self._lock = threading.Lock()

    @property
    def enabled(self) -> bool:
        return bool(self.path)

    @contextmanager
    def span(self, name: str, cat: str = 'step', **args):
        if not self.enabled:
            yield
            return
        t0 = time.perf_counter_ns() // 1000
        try:
            yield
        finally:
            t1 = time.perf_counter_ns() // 1000
            with self._lock:
                self.events.append({'name': name, 'cat': cat, 'ph': 'X', 'ts': t0, 'dur': t1 - t0,
                                    'pid': self.rank, 'tid': threading.get_ident() % 100000, 'args': args})

    def instant(self, name: str, **args):
        if self.enabled:
            with self._lock:
                self.events.append({'name': name, 'ph': 'i', 'ts': time.perf_counter_ns() // 1000,
                                    'pid': self.rank, 'tid': 0, 's': 'p', 'args': args})

    def dump(self) -> Optional[str]:
        if not self.enabled:
            return None
        path = self.path if self.rank == 0 else f'{self.path}.r{self.rank}'
        with open(path, 'w') as f:
            json.dump({'traceEvents': self.events, 'displayTimeUnit': 'ms'}, f)
        return path


_CLOCK_QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
                'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
                'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')


class ClockSampler:
    """Samples ``nvidia-smi`` clocks/throttle reasons every ``period_ms`` while running."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._lines: List[str] = []
        self._thread: Optional[threading.Thread] = None

    def start(self):
        try:
            self._proc = subprocess.Popen(
                ['nvidia-smi', f'--query-gpu={_CLOCK_QUERY}', '--format=csv,noheader,nounits',
                 '-i', str(self.gpu_index), '-lms', str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except (FileNotFoundError, OSError):
            self._proc = None
            return self

        def pump():
            for line in self._proc.stdout:
                self._lines.append(line.strip())
        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> dict:
        if self._proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable'], 'samples': 0}
        time.sleep(min(0.25, self.period_ms / 1000.0 * 2))
        self._proc.terminate()
        try:
            self._proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        sm, mx, power = [], [], []
        reasons = set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for ln in self._lines:
            parts = [p.strip() for p in ln.split(',')]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2])); power.append(float(parts[3]))
            except ValueError:
                continue
            for n, v in zip(names, parts[5:9]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['no samples'], 'samples': 0}
        # "under load" = samples at or above the median power draw
        order = sorted(range(len(sm)), key=lambda i: power[i])
        loaded = [sm[i] for i in order[len(order) // 2:]]
        loaded.sort()
        return {'sm_mhz': loaded[len(loaded) // 2], 'sm_max_mhz': max(mx), 'reasons': sorted(reasons),
                'samples': len(sm), 'power_w_max': max(power)}


_L2_FLUSH_BUF = {}


def flush_l2(device=None, nbytes: int = 256 << 20) -> None:
    """Evict L2 (126 MB on B200) by overwriting a 256 MB scratch buffer."""
    if not torch.cuda.is_available():
        return
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    buf = _L2_FLUSH_BUF.get(dev)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        _L2_FLUSH_BUF[dev] = buf
    buf.zero_()
