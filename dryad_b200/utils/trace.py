"""Device-side timeline of a job: NVTX ranges + CUDA-event spans per vertex / fused stage.

The reference's trace is the GM event log plus per-attempt wall-clock statistics
(GraphManager/vertex/DrVertexRecord.h:30-118, GraphManager/reporting/DrCalypsoReporting.cpp).  On a
GPU the host clock around an asynchronous launch measures nothing, so every span is a pair of CUDA
events recorded on the launching stream; they are resolved ONCE at job end (a single synchronize),
never inside the job.  NVTX ranges with the same names make the spans visible to Nsight tools.
Enabled by `DRYAD_TRACE=1` or `context.trace_gpu = True`; off by default (two event records per
vertex are cheap, but not free).
"""
from __future__ import annotations

import contextlib
import json
import os
from typing import Dict, List, Optional, Tuple


def enabled(context=None) -> bool:
    if context is not None and getattr(context, "trace_gpu", False):
        return True
    return os.environ.get("DRYAD_TRACE", "0") == "1"


class GpuTimeline:
    def __init__(self, device=None):
        self.device = device
        self._spans: List[Tuple[str, dict, object, object]] = []
        self._cuda = None

    def _has_cuda(self) -> bool:
        if self._cuda is None:
            try:
                import torch
                self._cuda = torch.cuda.is_available()
            except Exception:  # noqa: BLE001
                self._cuda = False
        return self._cuda

    @contextlib.contextmanager
    def span(self, name: str, **attrs):
        """Times the GPU work enqueued inside the block on the current stream."""
        if not self._has_cuda():
            yield
            return
        import torch
        torch.cuda.nvtx.range_push(name)
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        try:
            yield
        finally:
            e1.record()
            torch.cuda.nvtx.range_pop()
            self._spans.append((name, attrs, e0, e1))

    def resolve(self) -> List[Dict]:
        """One synchronize, then device milliseconds of every span in launch order."""
        if not self._spans:
            return []
        import torch
        torch.cuda.synchronize()
        out = []
        for name, attrs, e0, e1 in self._spans:
            d = {"name": name, "gpu_ms": round(e0.elapsed_time(e1), 4)}
            d.update(attrs)
            out.append(d)
        self._spans = []
        return out

    def write(self, path: str, spans: Optional[List[Dict]] = None) -> str:
        spans = self.resolve() if spans is None else spans
        with open(path, "w") as f:
            for s in spans:
                f.write(json.dumps(s) + "\n")
        return path
