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


class StrColumn:
    """A variable-length (string) column: `n` views (start, length) into one byte buffer, plus the
    64-bit key hash of every string when known (FNV-1a + mixer, equal on host and device).

    This is the offsets + bytes layout generalised to non-contiguous views, so a tokenizer can emit
    a column of words without copying a byte, and a GroupBy can hand out its representatives as
    views of the same buffer.  Reference: strings travel as length-prefixed UTF-8
    (LinqToDryad/DryadLinqBinaryWriter.cs:523-546); LineRecord (LinqToDryad/LineRecord.cs:30-176)."""

    def __init__(self, buf: torch.Tensor, starts: Optional[torch.Tensor], lens: Optional[torch.Tensor],
                 hashes: Optional[torch.Tensor] = None, whole_lines: bool = False):
        self.buf = buf                       # uint8 [nbytes]
        self._starts, self._lens = starts, lens
        self.hashes = hashes                 # int64 [n] or None
        self.whole_lines = whole_lines       # rows are exactly the '\n'-separated lines of buf, in order
        self.dtype = "str"

    # lines of a text buffer are located lazily: a SelectMany(split) right behind the source never
    # needs them (it tokenizes the buffer directly)
    def _materialise(self) -> None:
        if self._starts is None:
            from ..ops import core as K
            if self.buf.device.type == "cuda":
                self._starts, self._lens, self.hashes = K.tokenize(self.buf, "\n", True)
            else:
                data = bytes(self.buf.numpy())
                st, ln, pos = [], [], 0
                from ..oracle.text import split_lines
                for l in data.split(b"\n")[:-1] if data.endswith(b"\n") else data.split(b"\n"):
                    st.append(pos); ln.append(len(l)); pos += len(l) + 1
                self._starts = torch.tensor(st, dtype=torch.int64)
                self._lens = torch.tensor(ln, dtype=torch.int32)

    @property
    def starts(self) -> torch.Tensor:
        self._materialise()
        return self._starts

    @property
    def lens(self) -> torch.Tensor:
        self._materialise()
        return self._lens

    @property
    def shape(self):
        return (int(self.starts.numel()),)

    @property
    def device(self):
        return self.buf.device

    def dim(self) -> int:
        return 1

    def numel(self) -> int:
        return self.shape[0]

    def element_size(self) -> int:
        return 1

    def nbytes(self) -> int:
        return int(self.buf.numel()) + (0 if self._starts is None else 12 * int(self._starts.numel()))

    def key_hashes(self) -> torch.Tensor:
        """int64 hash per string (computed on first use)."""
        if self.hashes is None:
            from .. import _native as N
            from ..ops import core as K
            n = self.shape[0]
            out = torch.empty(n, dtype=torch.int64, device=self.device)
            if self.device.type == "cuda":
                N.check(N.kernels().dryad_hash_views(K.ptr(self.buf), K.ptr(self.starts.contiguous()), K.ptr(self.lens.contiguous()),
                                                     n, K.ptr(out), K.stream()), "hash_views")
                K._count()
            else:
                from ..utils.hashing import hash_bytes_py as hash_bytes_host
                data = bytes(self.buf.numpy())
                vals = [hash_bytes_host(data[a:a + b]) for a, b in zip(self.starts.tolist(), self.lens.tolist())]
                out = torch.tensor([v - (1 << 64) if v >= (1 << 63) else v for v in vals], dtype=torch.int64)
            self.hashes = out
        return self.hashes

    def to(self, device, non_blocking: bool = False) -> "StrColumn":
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)   # noqa: E731
        return StrColumn(mv(self.buf), mv(self._starts), mv(self._lens), mv(self.hashes), self.whole_lines)

    def index(self, idx: torch.Tensor) -> "StrColumn":
        """Row gather: the views are gathered, the bytes stay where they are."""
        i = idx.long()
        return StrColumn(self.buf, self.starts[i], self.lens[i], None if self.hashes is None else self.hashes[i])

    def slice(self, a: int, b: int) -> "StrColumn":
        return StrColumn(self.buf, self.starts[a:b], self.lens[a:b], None if self.hashes is None else self.hashes[a:b])

    @staticmethod
    def concat(cols: Sequence["StrColumn"]) -> "StrColumn":
        if len(cols) == 1:
            return cols[0]
        if all(c.buf.data_ptr() == cols[0].buf.data_ptr() and c.buf.numel() == cols[0].buf.numel() for c in cols):
            buf, shifts = cols[0].buf, [0] * len(cols)
        else:
            buf = torch.cat([c.buf for c in cols])
            shifts, acc = [], 0
            for c in cols:
                shifts.append(acc)
                acc += int(c.buf.numel())
        hs = None if any(c.hashes is None for c in cols) else torch.cat([c.hashes for c in cols])
        return StrColumn(buf, torch.cat([c.starts + s for c, s in zip(cols, shifts)]), torch.cat([c.lens for c in cols]), hs)

    def pack(self, width: int) -> torch.Tensor:
        """uint8 [n, width]: every string copied into a zero-padded fixed-width slot (for the
        fixed-width fused shuffle).  The caller guarantees lens <= width, width % 4 == 0."""
        from .. import _native as N
        from ..ops import core as K
        n = self.shape[0]
        out = torch.empty((n, width), dtype=torch.uint8, device=self.device)
        N.check(N.kernels().dryad_str_pack(K.ptr(self.buf), K.ptr(self.starts.contiguous()), K.ptr(self.lens.contiguous()), n, width,
                                           K.ptr(out), K.stream()), "str_pack")
        K._count()
        return out

    @staticmethod
    def from_packed(rows: torch.Tensor, lens: torch.Tensor, hashes: Optional[torch.Tensor] = None) -> "StrColumn":
        n, w = rows.shape
        starts = torch.arange(n, dtype=torch.int64, device=rows.device) * w
        return StrColumn(rows.reshape(-1), starts, lens.to(torch.int32), hashes)

    def to_list(self) -> List[str]:
        data = bytes(self.buf.cpu().numpy())
        return [data[a:a + b].decode("utf-8") for a, b in zip(self.starts.cpu().tolist(), self.lens.cpu().tolist())]


class ColumnBatch:
    def __init__(self, cols: Dict[str, torch.Tensor], n: Optional[int] = None, tuple_like: bool = False,
                 key_spec: Optional[tuple] = None, acc_lanes: Optional[int] = None):
        self.acc_lanes = acc_lanes          # partial-aggregate layout: records are (key, [acc...])
        self.cols = cols
        self._n = None if n is None else int(n)
        self.tuple_like = tuple_like
        self.key_spec = key_spec            # (offset, length) of the byte key in a `row` batch

    @property
    def n(self) -> int:
        if self._n is None:                 # lazily: a text column locates its lines only when asked
            self._n = int(next(iter(self.cols.values())).shape[0]) if self.cols else 0
        return self._n

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
        return {k: ("s" if isinstance(v, StrColumn) else _kind(v.dtype)) for k, v in self.cols.items()}

    @property
    def has_strings(self) -> bool:
        return any(isinstance(v, StrColumn) for v in self.cols.values())

    @property
    def device(self):
        return next(iter(self.cols.values())).device if self.cols else torch.device("cpu")

    def __len__(self):
        return self.n

    def nbytes(self) -> int:
        return sum(v.nbytes() if isinstance(v, StrColumn) else v.numel() * v.element_size() for v in self.cols.values())

    # ---- movement / slicing --------------------------------------------------------------------
    def to(self, device, non_blocking: bool = False) -> "ColumnBatch":
        return ColumnBatch({k: v.to(device, non_blocking=non_blocking) for k, v in self.cols.items()}, self.n,
                           self.tuple_like, self.key_spec, self.acc_lanes)

    def slice(self, a: int, b: int) -> "ColumnBatch":
        return ColumnBatch({k: (v.slice(a, b) if isinstance(v, StrColumn) else v[a:b]) for k, v in self.cols.items()},
                           max(0, min(b, self.n) - a), self.tuple_like, self.key_spec)

    def split(self, n: int) -> List["ColumnBatch"]:
        per = (self.n + n - 1) // n if self.n else 0
        return [self.slice(i * per, min(self.n, (i + 1) * per)) for i in range(n)]

    def take(self, idx: torch.Tensor) -> "ColumnBatch":
        """Row gather by int32 indices with the hand-written gather kernel on CUDA."""
        if self.device.type == "cuda":
            from ..ops import core as K
            out = {}
            for k, v in self.cols.items():
                if isinstance(v, StrColumn):
                    out[k] = v.index(idx)
                elif v.dim() == 1 and v.element_size() in (4, 8):
                    out[k] = K.gather_rows(v.contiguous(), idx)
                elif v.dim() == 2 and (v.shape[1] * v.element_size()) % 4 == 0:
                    out[k] = K.gather_rows(v.contiguous(), idx)
                else:
                    out[k] = v[idx.long()]
            return ColumnBatch(out, idx.numel(), self.tuple_like, self.key_spec, self.acc_lanes)
        return ColumnBatch({k: (v.index(idx) if isinstance(v, StrColumn) else v[idx.long()]) for k, v in self.cols.items()},
                           idx.numel(), self.tuple_like, self.key_spec, self.acc_lanes)

    @staticmethod
    def concat(batches: Sequence["ColumnBatch"]) -> "ColumnBatch":
        batches = [b for b in batches if b is not None]
        if len(batches) > 1:
            full = [b for b in batches if b.n > 0]         # empty pieces add nothing, and may lack the columns of the others
            batches = full or batches[:1]
        if len(batches) == 1:
            return batches[0]
        names = list(batches[0].cols)
        def cat(k):
            vs = [b.cols[k] for b in batches]
            return StrColumn.concat(vs) if isinstance(vs[0], StrColumn) else torch.cat(vs)
        return ColumnBatch({k: cat(k) for k in names},
                           sum(b.n for b in batches), batches[0].tuple_like, batches[0].key_spec, batches[0].acc_lanes)

    # ---- records (storage boundary / host path) --------------------------------------------
    def to_records(self) -> List:
        if self.n == 0:
            return []
        if self.is_rows:
            a = self.cols[ROW_COL].cpu().numpy()
            return [bytes(r) for r in a]
        host = {k: (v.to_list() if isinstance(v, StrColumn) else v.cpu().tolist()) for k, v in self.cols.items()}
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
