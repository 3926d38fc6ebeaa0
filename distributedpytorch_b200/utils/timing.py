"""Device-side timing (the reference only has host wall-clock).  CUDA events on the launching
stream, synchronise on both sides, max over ranks for anything multi-GPU."""
from __future__ import annotations

import time
from typing import Callable, List, Optional

import torch


class DeviceTimer:
    """Times a region with CUDA events (falls back to perf_counter on CPU)."""

    def __init__(self, device: Optional[torch.device] = None):
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self.ms = 0.0

    def __enter__(self):
        if self.cuda:
            torch.cuda.synchronize()
            self._e0 = torch.cuda.Event(enable_timing=True)
            self._e1 = torch.cuda.Event(enable_timing=True)
            self._e0.record()
        else:
            self._t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.cuda:
            self._e1.record()
            torch.cuda.synchronize()
            self.ms = self._e0.elapsed_time(self._e1)
        else:
            self.ms = (time.perf_counter() - self._t0) * 1e3
        return False


def max_over_ranks(value_ms: float) -> float:
    """All-reduce(max) of a scalar over the default process group, if one is initialised."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return value_ms
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor([value_ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def time_steps(step: Callable[[], None], steps: int, warmup: int) -> List[float]:
    for _ in range(warmup):
        step()
    out = []
    for _ in range(steps):
        with DeviceTimer() as t:
            step()
        out.append(t.ms)
    return out
