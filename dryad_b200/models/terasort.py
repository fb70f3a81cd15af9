"""TeraSort-style workload: 100-byte records, 10-byte big-endian key, RangePartition + OrderBy.

BASELINE.json config 2.  Data is synthetic (gensort-shaped: uniform random key bytes, arbitrary
payload) because there is no network; generation is deterministic per (seed, rank).
"""
from __future__ import annotations

import torch

RECORD_BYTES = 100
KEY_BYTES = 10


def generate_rows(n: int, seed: int, device="cuda", rank: int = 0) -> torch.Tensor:
    """[n, 100] uint8 records with uniform random bytes (key = first 10 bytes)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed * 1000003 + rank)
    words = RECORD_BYTES // 4
    out = torch.empty((n, words), dtype=torch.int32, device=device)
    chunk = 1 << 24
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        out[s:e] = torch.randint(-2**31, 2**31 - 1, (e - s, words), dtype=torch.int32, device=device, generator=g)
    return out.view(torch.uint8).view(n, RECORD_BYTES)


def key_checksum(rows: torch.Tensor) -> int:
    """Wrapping sum of the 64-bit key prefixes (permutation invariant), computed with torch."""
    k = rows[:, :8].to(torch.int64)
    v = torch.zeros(rows.shape[0], dtype=torch.int64, device=rows.device)
    for b in range(8):
        v = (v << 8) | k[:, b]
    return int(v.sum().item()) & ((1 << 64) - 1)


def validate_sorted(rows_cpu: torch.Tensor) -> bool:
    import numpy as np
    a = rows_cpu[:, :KEY_BYTES].numpy()
    keys = [bytes(r) for r in a]
    return all(x <= y for x, y in zip(keys, keys[1:]))


def result_rows(result_query) -> torch.Tensor:
    """The sorted rows this process holds after `order_by(...).submit_and_wait()` as ONE device
    tensor [m, RECORD_BYTES] (the partitions of a row table are views of the sorter's output)."""
    t = result_query.expr.params["table"]
    parts = [p.cols["row"] for p in t.partitions if p is not None and len(p)]
    if not parts:
        return torch.empty((0, RECORD_BYTES), dtype=torch.uint8)
    return parts[0] if len(parts) == 1 else torch.cat(parts)


def validate(result_query) -> list:
    """TeraValidate-style summary of a sorted result held by this process: [order violations,
    wrapping key checksum, row count].  Reads 24 bytes back from the device."""
    from ..ops import core as K
    t = result_query.expr.params["table"]
    total = None
    rows_n = 0
    for p in t.partitions:
        if p is None or len(p) == 0:
            continue
        rows = p.cols["row"]
        off, ln = p.key_spec or (0, KEY_BYTES)
        r = K.check_sorted_rows(rows.contiguous(), off, ln)
        total = r if total is None else total + r
        rows_n += rows.shape[0]
    if total is None:
        return [0, 0, 0]
    host = total.cpu()
    return [int(host[0]), int(host[1]) & ((1 << 64) - 1), rows_n]
