"""Columnar partitions: struct-of-arrays of torch tensors (HBM or host), the in-engine data model.

Records exist only at the storage boundary and on the host path; inside the GPU engine a
partition is a `ColumnBatch`: name -> 1-D tensor per fixed-width column, or a single 2-D uint8
`row` column for fixed-width byte records (TeraSort-style).  Schema kinds follow the expression
IR ('i' int64, 'f' float64, 'b' bool); 'rows' marks a byte-record batch.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch

from ..api.types import Record

ROW_COL = "row"


def _kind(dt: torch.dtype) -> str:
    if dt in (torch.float32, torch.float64):
        return "f"
    if dt == torch.bool:
        return "b"
    return "i"


class ColumnBatch:
    def __init__(self, cols: Dict[str, torch.Tensor], n: Optional[int] = None, tuple_like: bool = False,
                 key_spec: Optional[tuple] = None, acc_lanes: Optional[int] = None):
        self.acc_lanes = acc_lanes          # partial-aggregate layout: records are (key, [acc...])
        self.cols = cols
        self.n = int(n if n is not None else (next(iter(cols.values())).shape[0] if cols else 0))
        self.tuple_like = tuple_like
        self.key_spec = key_spec            # (offset, length) of the byte key in a `row` batch

    # ---- construction ------------------------------------------------------------------------
    @staticmethod
    def from_columns(columns: Dict[str, Any], device=None) -> "ColumnBatch":
        cols = {}
        for k, v in columns.items():
            t = v if isinstance(v, torch.Tensor) else torch.as_tensor(np.asarray(v))
            if t.dtype in (torch.int32, torch.int16, torch.int8, torch.uint8) and t.dim() == 1:
                t = t.to(torch.int64)
            if t.dtype == torch.float32 and t.dim() == 1:
                t = t.to(torch.float64)
            cols[k] = t.to(device) if device is not None else t
        return ColumnBatch(cols)

    @staticmethod
    def from_records(recs: Sequence, schema: Dict[str, str], device=None) -> "ColumnBatch":
        dt = {"i": torch.int64, "f": torch.float64, "b": torch.bool}
        names = list(schema)
        if names == ["value"]:
            t = torch.tensor(list(recs), dtype=dt[schema["value"]])
            cols = {"value": t}
        else:
            cols = {}
            for j, k in enumerate(names):
                vals = [(r[k] if isinstance(r, dict) else (getattr(r, k) if hasattr(r, k) else r[j])) for r in recs]
                cols[k] = torch.tensor(vals, dtype=dt[schema[k]])
        if device is not None:
            cols = {k: v.to(device) for k, v in cols.items()}
        return ColumnBatch(cols, len(recs))

    @staticmethod
    def rows(t: torch.Tensor, key_off: int = 0, key_len: int = 8) -> "ColumnBatch":
        assert t.dim() == 2 and t.dtype == torch.uint8
        return ColumnBatch({ROW_COL: t}, t.shape[0], key_spec=(key_off, key_len))

    # ---- properties --------------------------------------------------------------------------
    @property
    def is_rows(self) -> bool:
        return ROW_COL in self.cols and self.cols[ROW_COL].dim() == 2

    @property
    def schema(self) -> Dict[str, str]:
        if self.is_rows:
            return {ROW_COL: "o"}
        return {k: _kind(v.dtype) for k, v in self.cols.items()}

    @property
    def device(self):
        return next(iter(self.cols.values())).device if self.cols else torch.device("cpu")

    def __len__(self):
        return self.n

    def nbytes(self) -> int:
        return sum(v.numel() * v.element_size() for v in self.cols.values())

    # ---- movement / slicing --------------------------------------------------------------------
    def to(self, device, non_blocking: bool = False) -> "ColumnBatch":
        return ColumnBatch({k: v.to(device, non_blocking=non_blocking) for k, v in self.cols.items()}, self.n,
                           self.tuple_like, self.key_spec, self.acc_lanes)

    def slice(self, a: int, b: int) -> "ColumnBatch":
        return ColumnBatch({k: v[a:b] for k, v in self.cols.items()}, max(0, min(b, self.n) - a), self.tuple_like,
                           self.key_spec)

    def split(self, n: int) -> List["ColumnBatch"]:
        per = (self.n + n - 1) // n if self.n else 0
        return [self.slice(i * per, min(self.n, (i + 1) * per)) for i in range(n)]

    def take(self, idx: torch.Tensor) -> "ColumnBatch":
        """Row gather by int32 indices with the hand-written gather kernel on CUDA."""
        if self.device.type == "cuda":
            from ..ops import core as K
            out = {}
            for k, v in self.cols.items():
                if v.dim() == 1 and v.element_size() in (4, 8):
                    out[k] = K.gather_rows(v.contiguous(), idx)
                elif v.dim() == 2 and (v.shape[1] * v.element_size()) % 4 == 0:
                    out[k] = K.gather_rows(v.contiguous(), idx)
                else:
                    out[k] = v[idx.long()]
            return ColumnBatch(out, idx.numel(), self.tuple_like, self.key_spec, self.acc_lanes)
        return ColumnBatch({k: v[idx.long()] for k, v in self.cols.items()}, idx.numel(), self.tuple_like, self.key_spec, self.acc_lanes)

    @staticmethod
    def concat(batches: Sequence["ColumnBatch"]) -> "ColumnBatch":
        batches = [b for b in batches if b is not None]
        if len(batches) == 1:
            return batches[0]
        names = list(batches[0].cols)
        return ColumnBatch({k: torch.cat([b.cols[k] for b in batches]) for k in names},
                           sum(b.n for b in batches), batches[0].tuple_like, batches[0].key_spec, batches[0].acc_lanes)

    # ---- records (storage boundary / host path) --------------------------------------------
    def to_records(self) -> List:
        if self.n == 0:
            return []
        if self.is_rows:
            a = self.cols[ROW_COL].cpu().numpy()
            return [bytes(r) for r in a]
        host = {k: v.cpu().tolist() for k, v in self.cols.items()}
        names = list(host)
        if self.acc_lanes is not None:
            lanes = [host[f"agg{i}"] for i in range(self.acc_lanes)]
            return [(k, [l[j] for l in lanes]) for j, k in enumerate(host["key"])]
        if names == ["value"]:
            return host["value"]
        if self.tuple_like:
            cls = self.tuple_like if isinstance(self.tuple_like, type) else tuple
            rows = zip(*[host[k] for k in names])
            return list(rows) if cls is tuple else [cls(*r) for r in rows]
        return [Record(zip(names, vals)) for vals in zip(*[host[k] for k in names])]
