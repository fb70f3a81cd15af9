"""Execution backends: where a vertex's operator code runs.

`HostBackend` — Python records on the CPU (general path; LocalJobSubmission-style multi-partition
execution in one process).  `GpuBackend` — columnar partitions in HBM with the sm_100a kernels for
every stage whose lambdas traced; anything else drops to the host path for that vertex (the
documented slow path) and comes back columnar when the schema allows.
"""
from __future__ import annotations

from . import vertex_ops


class HostBackend:
    name = "host"
    wants_records = True

    def __init__(self, context):
        self.context = context

    def run_vertex(self, node, part, inputs, env):
        # raw text partitions stay raw: a native vertex operator may consume the bytes directly, and
        # they turn into LineRecords only when a vertex iterates them
        ins = [[p if getattr(p, "is_text_partition", False) else self.to_records(p) for p in payloads] for payloads in inputs]
        return vertex_ops.run(node, part, ins, env)

    @staticmethod
    def count(payload) -> int:
        if getattr(payload, "is_text_partition", False) and getattr(payload, "_lines", 1) is None:
            return payload.line_count()              # raw text: count lines without materialising them
        try:
            return len(payload)
        except TypeError:
            return 0

    @staticmethod
    def to_records(payload):
        return payload.to_records() if hasattr(payload, "to_records") else payload


_cache = {}


def get_backend(context):
    key = (id(context), context.device)
    b = _cache.get(key)
    if b is None:
        if str(context.device).startswith("cuda"):
            from .gpu_backend import GpuBackend
            b = GpuBackend(context)
        else:
            b = HostBackend(context)
        _cache[key] = b
    return b
