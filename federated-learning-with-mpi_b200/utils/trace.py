"""Optional tracing: NVTX ranges around the phases of a round (visible in Nsight Systems / ncu's NVTX
filters) and a tiny wall-clock phase log.  The reference has no tracing at all; enable with
``FEDB200_NVTX=1`` or ``trace.enable()``.  Disabled, the context manager costs one attribute lookup."""
from __future__ import annotations

import contextlib
import os
import time
from typing import Dict, List

_enabled = os.environ.get("FEDB200_NVTX") is not None
phase_log: List[Dict] = []


def enable(on: bool = True) -> None:
    global _enabled
    _enabled = on


@contextlib.contextmanager
def span(name: str):
    if not _enabled:
        yield
        return
    import torch
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.nvtx.range_push(name)
    t0 = time.perf_counter()
    try:
        yield
    finally:
        if cuda:
            torch.cuda.nvtx.range_pop()
        phase_log.append({"phase": name, "host_ms": (time.perf_counter() - t0) * 1e3})
