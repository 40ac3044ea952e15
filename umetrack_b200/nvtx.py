"""NVTX ranges around the stages of the path (crop cameras, resample, encoder, time scan, FK + metrics),
for timelines taken with Nsight Systems.  Off unless UMETRACK_NVTX=1: a range is two calls into the
NVTX library per stage, which the per-frame tracker (under 1 ms per frame) should not pay by default."""
from __future__ import annotations

import contextlib
import os

_ON = os.environ.get("UMETRACK_NVTX", "0") not in ("", "0")
_NULL = contextlib.nullcontext()


def enabled() -> bool:
    return _ON


def enable(on: bool = True) -> None:
    global _ON
    _ON = bool(on)


def stage(name: str):
    """Context manager: an NVTX range `umetrack/<name>` when enabled, else nothing."""
    if not _ON:
        return _NULL
    import torch

    return torch.cuda.nvtx.range("umetrack/" + name)


# UMETRACK_STAGE_TIMES=1 (tools/profile_batched.py): push / pop also synchronise the device and add the
# stage's wall time to `stage_times` -- a serialised breakdown, not a measurement of the pipelined run.
_TIMES = os.environ.get("UMETRACK_STAGE_TIMES", "0") not in ("", "0")
stage_times: dict = {}
_open: list = []


def push(name: str) -> None:
    """Open `umetrack/<name>` (for stages that are sections of a long function); close with pop()."""
    if _ON:
        import torch

        torch.cuda.nvtx.range_push("umetrack/" + name)
    if _TIMES:
        import time

        import torch

        torch.cuda.synchronize()
        _open.append((name, time.perf_counter()))


def pop() -> None:
    if _ON:
        import torch

        torch.cuda.nvtx.range_pop()
    if _TIMES and _open:
        import time

        import torch

        torch.cuda.synchronize()
        name, t0 = _open.pop()
        stage_times[name] = stage_times.get(name, 0.0) + time.perf_counter() - t0
