"""Sub-graph vertex = CUDA graph.

The reference can run a whole connected sub-graph of vertices inside ONE process as a single
schedulable unit (DryadVertex/VertexHost/system/dprocess/src/subgraphvertex.cpp: the internal edges
become in-process FIFOs and the GM sees one vertex).  On a GPU the analogue of "many small operators
scheduled as one unit" is a CUDA graph: a chain of kernel launches whose shapes and buffers do not
change between executions is captured once and replayed with a single launch, which removes the
per-kernel launch latency from launch-bound inner loops (iterative jobs, small partitions).

Only launch sequences WITHOUT host round-trips can be captured: no `.item()`, no size-dependent
allocation, no epoch-numbered cross-GPU flags baked into kernel parameters.  `GraphVertex` checks
nothing of this itself — a capture that syncs raises inside CUDA; callers use it for the loops that
qualify (single-GPU k-means iterations, fused expression pipelines on fixed-size partitions).
"""
from __future__ import annotations

from typing import Any, Callable, Optional

import torch


class GraphVertex:
    def __init__(self, body: Callable[[], Any], warmup: int = 2, device=None):
        """`body` enqueues GPU work on the current stream using only pre-allocated tensors."""
        self.body = body
        self.device = device
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.result: Any = None
        self.replays = 0
        self._warmup = warmup

    def capture(self) -> "GraphVertex":
        # run eagerly on a side stream first (lazy module loading, attribute setting, allocator warm-up)
        s = torch.cuda.Stream(device=self.device)
        s.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(s):
            for _ in range(self._warmup):
                self.body()
        torch.cuda.current_stream(self.device).wait_stream(s)
        torch.cuda.synchronize(self.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.result = self.body()
        self.graph = g
        return self

    def run(self) -> Any:
        """Replay (captures on first use).  Outputs are the tensors `body` wrote during capture."""
        if self.graph is None:
            self.capture()
        self.graph.replay()
        self.replays += 1
        return self.result
