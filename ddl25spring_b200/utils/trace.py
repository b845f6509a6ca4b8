"""Tracing, metrics logging and hang detection.

The reference has ``perf_counter`` spans inside ``run()`` (hfl_complete.py:208-211,274-307), prints to
redirected stdout files and tqdm bars; a hung rank hangs everyone forever (homework-1 cell 48).

* :func:`span`       – NVTX range (shows up in nsys/ncu timelines) + host wall time + optional CUDA-event
                       device time, accumulated in a process-wide :class:`Tracer`.
* :class:`JsonlLogger` – one JSON object per line (round metrics, losses), rank-tagged.
* :class:`Watchdog`  – a daemon thread that aborts the process with a diagnostic when no progress
                       heartbeat arrives for ``timeout_s``: a deadlocked collective becomes a failure.
"""
from __future__ import annotations

import contextlib
import json
import os
import sys
import threading
import time
from collections import defaultdict

import torch


class Tracer:
    def __init__(self):
        self.host_s = defaultdict(float)
        self.device_ms = defaultdict(float)
        self.count = defaultdict(int)
        self._pending: list = []

    def flush(self) -> None:
        for name, e0, e1 in self._pending:
            e1.synchronize()
            self.device_ms[name] += e0.elapsed_time(e1)
        self._pending = []

    def report(self) -> dict:
        self.flush()
        return {k: {"calls": self.count[k], "host_s": round(self.host_s[k], 6), "device_ms": round(self.device_ms.get(k, 0.0), 4)}
                for k in self.count}


TRACER = Tracer()


@contextlib.contextmanager
def span(name: str, device_time: bool = False):
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.nvtx.range_push(name)
    ev = None
    if cuda and device_time:
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record()
    t0 = time.perf_counter()
    try:
        yield
    finally:
        TRACER.host_s[name] += time.perf_counter() - t0
        TRACER.count[name] += 1
        if ev is not None:
            ev[1].record()
            TRACER._pending.append((name, ev[0], ev[1]))
        if cuda:
            torch.cuda.nvtx.range_pop()


class JsonlLogger:
    def __init__(self, path: str | None = None, rank: int | None = None):
        self.rank = int(os.environ.get("RANK", "0")) if rank is None else rank
        self.f = open(path, "a") if path else None

    def log(self, **kv) -> None:
        rec = {"t": round(time.time(), 3), "rank": self.rank, **kv}
        line = json.dumps(rec)
        if self.f:
            self.f.write(line + "\n")
            self.f.flush()
        else:
            print(line, flush=True)


class Watchdog:
    """``wd = Watchdog(60); wd.start(); ... wd.beat() every iteration ...; wd.stop()``."""

    def __init__(self, timeout_s: float, on_timeout=None, what: str = "training loop"):
        self.timeout_s, self.what = timeout_s, what
        self._last = time.monotonic()
        self._stop = threading.Event()
        self._thr: threading.Thread | None = None
        self.on_timeout = on_timeout or self._abort
        self.fired = False

    def _abort(self):
        sys.stderr.write(f"[watchdog] no progress in {self.what} for {self.timeout_s}s on rank {os.environ.get('RANK', '0')}: "
                         "aborting so the launcher sees a failure instead of a hang\n")
        sys.stderr.flush()
        os._exit(124)

    def _loop(self):
        while not self._stop.wait(min(1.0, self.timeout_s / 4)):
            if time.monotonic() - self._last > self.timeout_s:
                self.fired = True
                self.on_timeout()
                return

    def start(self):
        self._last = time.monotonic()
        self._thr = threading.Thread(target=self._loop, daemon=True)
        self._thr.start()
        return self

    def beat(self):
        self._last = time.monotonic()

    def stop(self):
        self._stop.set()
