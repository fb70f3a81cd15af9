"""Python face of the native graph manager (csrc/host/sched.cpp): the job state machine with
versioned vertex attempts, failure budget, upstream invalidation, speculative duplicates and the
per-stage straggler model (reference: GraphManager/vertex/*, GraphManager/stagemanager/*)."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Tuple

from .. import _native as N

JOB_STATES = {0: "NotStarted", 1: "Running", 2: "Stopping", 3: "Succeeded", 4: "Failed"}
VERTEX_STATES = {0: "Waiting", 1: "Ready", 2: "Running", 3: "Completed", 4: "Failed", 5: "Removed"}
POINTWISE, CROSS_PRODUCT = 0, 1
CHANNEL_FILE, CHANNEL_FIFO = 0, 1


class GraphManager:
    def __init__(self, max_failures: int = 6, speculative: bool = True):
        self.lib = N.host()
        self.h = self.lib.gm_create(max_failures, int(speculative))
        self.stage_first_vertex: List[int] = []
        self._nverts = 0

    def close(self):
        if self.h:
            self.lib.gm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def add_stage(self, name: str, nparts: int) -> int:
        sid = self.lib.gm_add_stage(self.h, name.encode(), nparts)
        self.stage_first_vertex.append(self._nverts)
        self._nverts += nparts
        return sid

    def connect(self, src: int, dst: int, kind: int = POINTWISE, channel: int = CHANNEL_FILE) -> None:
        """channel=CHANNEL_FIFO streams the edge: producer and consumer vertices form a start clique
        (gang start / gang restart, GraphManager/vertex/DrClique.h:28-64)."""
        rc = self.lib.gm_connect_ex(self.h, src, dst, kind, channel)
        if rc:
            raise ValueError(f"cannot connect stage {src} -> {dst} (rc {rc})")

    def resize_stage(self, stage: int, new_parts: int, follow: bool = True) -> int:
        """Change the width of a stage that has not started (dynamic distributor); with `follow`
        the pointwise stages downstream of equal width are resized with it
        (DrPipelineSplitManager / DrSemiPipelineSplitManager).  Returns the number of stages resized."""
        rc = self.lib.gm_resize_stage(self.h, stage, new_parts, int(follow))
        if rc < 0:
            raise ValueError(f"stage {stage} cannot be resized to {new_parts} (rc {rc})")
        return rc

    def splice(self, stage: int, inputs: List[int], consumers: List[int]) -> int:
        """Run-time graph rewrite: ONE new vertex in `stage` (made with `add_stage(name, 0)`) reads
        `inputs`; in every consumer the edges from those inputs become one edge from the new vertex
        (DrDynamicAggregateManager internal vertices, DrDynamicBroadcastManager copy vertices).
        Returns the new vertex id."""
        a = (C.c_int * len(inputs))(*inputs)
        b = (C.c_int * len(consumers))(*consumers)
        v = self.lib.gm_splice(self.h, stage, a, len(inputs), b, len(consumers))
        if v < 0:
            raise ValueError(f"cannot splice a vertex into stage {stage} (rc {v})")
        self._nverts = max(self._nverts, v + 1)
        return v

    def vertex_inputs(self, vertex: int) -> List[int]:
        n = self.lib.gm_vertex_inputs(self.h, vertex, None, 0)
        out = (C.c_int * max(n, 1))()
        self.lib.gm_vertex_inputs(self.h, vertex, out, n)
        return list(out)[:n]

    def stage_vertices(self, stage: int) -> List[int]:
        n = self.lib.gm_stage_vertices(self.h, stage, None, 0)
        out = (C.c_int * max(n, 1))()
        self.lib.gm_stage_vertices(self.h, stage, out, n)
        return list(out)[:n]

    def clique(self, vertex: int) -> List[int]:
        """Members of the vertex's start clique ([vertex] when it starts alone)."""
        n = self.lib.gm_vertex_clique(self.h, vertex, None, 0)
        out = (C.c_int * max(n, 1))()
        self.lib.gm_vertex_clique(self.h, vertex, out, n)
        return list(out)[:n]

    @property
    def live_vertices(self) -> int:
        return self.lib.gm_live_vertex_count(self.h)

    @property
    def gang_restarts(self) -> int:
        return self.lib.gm_gang_restarts(self.h)

    def set_outlier_params(self, default_s: float = 600.0, min_s: float = 10.0, nonparam_fraction: float = 0.5):
        self.lib.gm_set_outlier_params(self.h, default_s, min_s, nonparam_fraction)

    def start(self) -> None:
        self.lib.gm_start(self.h)

    def next_ready(self, now: float) -> Optional[Tuple[int, int]]:
        v, ver = C.c_int(), C.c_int()
        rc = self.lib.gm_next_ready(self.h, now, C.byref(v), C.byref(ver))
        return (v.value, ver.value) if rc == 1 else None

    def completed(self, vertex: int, version: int, seconds: float, bytes_in: float = 0.0) -> bool:
        return self.lib.gm_report_completed(self.h, vertex, version, seconds, bytes_in) == 1

    def failed(self, vertex: int, version: int, message: str = "", bad_input: int = -1) -> int:
        """0 = will be retried, 1 = job failed, 2 = upstream producer invalidated."""
        return self.lib.gm_report_failed(self.h, vertex, version, bad_input, message.encode()[:2000])

    def check_duplicates(self, now: float) -> int:
        return self.lib.gm_check_duplicates(self.h, now)

    @property
    def state(self) -> str:
        return JOB_STATES[self.lib.gm_job_state(self.h)]

    def vertex(self, v: int) -> dict:
        a = [C.c_int() for _ in range(5)]
        self.lib.gm_vertex_info(self.h, v, *[C.byref(x) for x in a])
        return {"stage": a[0].value, "part": a[1].value, "state": VERTEX_STATES[a[2].value],
                "completed_version": a[3].value, "failures": a[4].value}

    def counters(self) -> dict:
        a = [C.c_int() for _ in range(4)]
        self.lib.gm_counters(self.h, *[C.byref(x) for x in a])
        return {"completed": a[0].value, "failures": a[1].value, "duplicates": a[2].value, "invalidations": a[3].value}

    def error(self) -> str:
        buf = C.create_string_buffer(4096)
        self.lib.gm_error(self.h, buf, 4096)
        return buf.value.decode(errors="replace")

    def stage_model(self, stage: int, bytes_: float = 0.0) -> dict:
        a = [C.c_double() for _ in range(4)]
        have = self.lib.gm_stage_model(self.h, stage, *[C.byref(x) for x in a], bytes_)
        return {"have_model": have == 1, "startup": a[0].value, "per_byte": a[1].value, "sigma": a[2].value,
                "outlier_threshold": a[3].value}


def aggregate_groups(bytes_: List[int], machine: List[int], pod: List[int], max_inputs: int = 128,
                     max_bytes: int = 1 << 30, skip_bytes: int = 512 << 20) -> List[int]:
    """Runtime aggregation-tree grouping (DrDynamicAggregateManager.h:174-179)."""
    n = len(bytes_)
    out = (C.c_int32 * n)()
    N.host().dyn_aggregate_groups(n, (C.c_int64 * n)(*bytes_), (C.c_int32 * n)(*machine), (C.c_int32 * n)(*pod),
                                  max_inputs, max_bytes, skip_bytes, out)
    return list(out)


def consumer_count(total_bytes: int, data_per_vertex: int = 1 << 30, lo: int = 1, hi: int = 20000) -> int:
    return N.host().dyn_consumer_count(total_bytes, data_per_vertex, lo, hi)


def range_consumer_count(sample_bytes: int, sampling_rate: float = 0.001, data_per_vertex: int = 1 << 30,
                         lo: int = 1, hi: int = 20000) -> int:
    """Consumers of a dynamic range distributor: data volume = sample volume / sampling rate
    (GraphManager/stagemanager/DrDynamicRangeDistributor.h:22-78)."""
    return N.host().dyn_range_consumer_count(sample_bytes, sampling_rate, data_per_vertex, lo, hi)


def broadcast_tree(consumers: int) -> Tuple[int, List[int]]:
    out = (C.c_int32 * max(consumers, 1))()
    copies = N.host().dyn_broadcast_tree(consumers, out)
    return copies, list(out)[:consumers]
