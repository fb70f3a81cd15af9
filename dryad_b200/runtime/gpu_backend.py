"""GPU vertex runtime: physical operators over ColumnBatch partitions with the sm_100a kernels.

Counterpart of the host operator library for stages whose lambdas traced to column expressions
(reference operator semantics: LinqToDryad/DryadLinqVertex.cs; kernel inventory SURVEY §2.15).
A vertex whose operator, lambdas or payloads are not GPU-expressible runs on the host path for
that vertex only (records), and its output is re-columnarised when the schema allows.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence

import torch

from ..api.errors import DryadLinqException, ErrorCode
from ..expr import ir as X
from ..expr.compile import CompileError, run_on_gpu
from ..ops import core as K
from ..plan.qnodes import DLinqQueryNode, NodeType
from . import vertex_ops
from .columnar import ROW_COL, ColumnBatch, StrColumn


def _tuple_cls(ir):
    """False for record / scalar results; else the tuple class the lambda built (tuple, Pair, ...)."""
    return ir.cls if isinstance(ir, X.TupleFields) else False


def _has_reserved_key(*key_bits: torch.Tensor) -> bool:
    """The hashed tables reserve the bit pattern ~0 (int64 -1, a NaN payload for floats) as their
    EMPTY marker.  A vertex whose keys contain it runs on the host path instead (one pass + one
    host sync per vertex; the direct-addressed dense-key table has no reserved value)."""
    return any(t.numel() and bool((t == -1).any().item()) for t in key_bits)


def _numeric_schema(schema) -> bool:
    return bool(schema) and all(k in ("i", "f", "b") for k in schema.values())


def _bind_leaves(e: X.Expr, agg_kinds: Sequence[str]) -> X.Expr:
    """Replace `agg` leaves by columns agg<i> and key references by key columns."""
    if e.op == "agg":
        i = e.args[0]
        return X.col(f"agg{i}", agg_kinds[i])
    if e.op in ("col", "lit", "rowid"):
        return e
    return X.Expr(e.op, tuple(_bind_leaves(a, agg_kinds) if isinstance(a, X.Expr) else a for a in e.args), e.kind)


class GpuBackend:
    name = "gpu"
    wants_records = False

    def __init__(self, context):
        self.context = context
        self.device = torch.device(context.device)
        K.N.kernels()                      # fail loudly when the CUDA library is missing
        self.host_fallbacks = 0
        self.fallback_log: List[str] = []
        self._arena = None

    # ---- plumbing ------------------------------------------------------------------------------
    @staticmethod
    def count(payload) -> int:
        try:
            return len(payload)
        except TypeError:
            return 0

    @staticmethod
    def to_records(payload):
        return payload.to_records() if hasattr(payload, "to_records") else payload

    def _columnar(self, payload, schema) -> Any:
        if isinstance(payload, ColumnBatch):
            return payload if payload.device.type == "cuda" else payload.to(self.device, non_blocking=True)
        if _numeric_schema(schema) and isinstance(payload, list):
            if payload and isinstance(payload[0], tuple) and len(schema) > 1:
                b = ColumnBatch.from_records(payload, schema, self.device)
                b.tuple_like = True
                return b
            try:
                return ColumnBatch.from_records(payload, schema, self.device)
            except Exception:  # noqa: BLE001
                return payload
        return payload

    def run_vertex(self, node: DLinqQueryNode, part: int, inputs, env):
        if node.type == NodeType.Super:
            cur = inputs
            out = None
            for m in node.members:
                out = self._run_one(m, part, cur, env)
                cur = [[out[0]]]
            return out
        return self._run_one(node, part, inputs, env)

    def _run_one(self, node, part, inputs, env):
        fn = _GPU_OPS.get(node.type)
        all_col = all(isinstance(p, ColumnBatch) for payloads in inputs for p in payloads)
        if all_col and node.type not in _STR_AWARE and any(p.has_strings for payloads in inputs for p in payloads):
            fn = None                      # operators without a string-column form run on the host path
        if fn is not None and (all_col or node.type == NodeType.InputTable):
            try:
                r = fn(self, node, part, inputs, env)
            except CompileError:
                r = NotImplemented
            if r is not NotImplemented:
                return r
        self.host_fallbacks += 1
        self.fallback_log.append(f"{node.type.value}#{node.uid}")     # which vertices ran on the host path
        ins = [[self.to_records(p) for p in payloads] for payloads in inputs]
        ports = vertex_ops.run(node, part, ins, env)
        if node.n_out_ports == 1 and node.type not in (NodeType.GroupBy,):
            return [self._columnar(ports[0], node.schema)]
        return ports

    def _cat(self, payloads) -> ColumnBatch:
        return ColumnBatch.concat(list(payloads))


def _numeric_cols(b: ColumnBatch) -> Dict[str, torch.Tensor]:
    return {k: v for k, v in b.cols.items() if not isinstance(v, StrColumn)}


def _str_col(src: ColumnBatch, lam) -> Optional[StrColumn]:
    """The string column a traced key / selector names directly (`w => w`, `r => r.word`), else None."""
    ir = getattr(lam, "ir", None)
    if isinstance(ir, X.Expr) and ir.op == "col" and ir.kind == "s":
        c = src.cols.get(ir.args[0])
        return c if isinstance(c, StrColumn) else None
    return None


# ---- operators -----------------------------------------------------------------------------------
def _input(be: GpuBackend, node, part, inputs, env):
    p = node.params["table"].partitions[part]
    data = getattr(p, "data", None)
    if isinstance(data, (bytes, bytearray)):
        # text partition: the bytes go to HBM as they are; the string column is views into them
        # (LinqToDryad/DryadLinqTextReader.cs:36-280 decodes lines on the CPU instead)
        dbuf = getattr(p, "device_buf", None)
        if dbuf is not None:
            dbuf = dbuf if dbuf.device.type == "cuda" else dbuf.to(be.device, non_blocking=True)
            if dbuf.numel() and bool((dbuf == 13).any().item()):
                return NotImplemented
            return [ColumnBatch({"value": StrColumn(dbuf, None, None, None, whole_lines=True)})]
        if b"\r" in data:
            return NotImplemented          # CRLF text: the host reader strips the '\r'
        if len(data) == 0:
            return [ColumnBatch({"value": StrColumn(torch.empty(0, dtype=torch.uint8, device=be.device),
                                                   torch.empty(0, dtype=torch.int64, device=be.device),
                                                   torch.empty(0, dtype=torch.int32, device=be.device))}, 0)]
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")                # read-only buffer: it is only copied from
            host = torch.frombuffer(data, dtype=torch.uint8)
        return [ColumnBatch({"value": StrColumn(host.to(be.device, non_blocking=True), None, None, None, whole_lines=True)})]
    return [be._columnar(p, node.schema)]


def _apply(be, node, part, inputs, env):
    """Apply / ApplyPerPartition with a DEVICE function: a callable marked `device = True` receives
    the ColumnBatches of its inputs as they are (tensors in HBM) and returns a ColumnBatch or a dict
    of tensors — user code that launches its own kernels, e.g. models.kmeans.KMeansStep.  Ordinary
    Python functions over records take the host path (LinqToDryad/DryadLinqVertex.cs:4726-4785)."""
    p = node.params
    fn = p["fn"].fn
    if not getattr(fn, "device", False) or p.get("with_index"):
        return NotImplemented
    args = [be._cat(payloads) for payloads in inputs]
    out = fn(*args)
    if isinstance(out, dict):
        out = ColumnBatch(out)
    return [out]


def _select_many(be, node, part, inputs, env):
    """SelectMany(l => l.Split(c)) over the lines of a text buffer = the tokenizer kernel (K13): the
    result is a string column of views into the SAME buffer, with the tokens' key hashes."""
    p = node.params
    ir = p["fn"].ir
    if p["indexed"] or p.get("result") is not None or not isinstance(ir, X.Expr) or ir.op != "split":
        return NotImplemented
    src = be._cat(inputs[0])
    a = ir.args[0]
    sc = src.cols.get(a.args[0]) if a.op == "col" else None
    if not isinstance(sc, StrColumn) or not sc.whole_lines or sc.device.type != "cuda":
        return NotImplemented
    if sc.buf.numel() == 0:
        return [src]
    starts, lens, hashes = K.tokenize(sc.buf, ir.args[1], True)
    return [ColumnBatch({"value": StrColumn(sc.buf, starts, lens, hashes)}, starts.numel())]


def _passthrough(be, node, part, inputs, env):
    sk = node.params.get("sort_key")
    if sk is not None:
        # ordered merge of key-sorted runs (MergeSort over the producer ports): merge-path k-way merge
        if node.params.get("comparer") is not None or not isinstance(sk.ir, X.Expr):
            return NotImplemented
        runs, base, batches = [], 0, [b for b in inputs[0] if b.n]
        if len(batches) <= 1:
            return [be._cat(inputs[0])]
        for b in batches:
            out, _ = run_on_gpu({"k": sk.ir}, b.cols, b.n)
            bits = K.sort_key_bits(out["k"], descending=bool(node.params.get("descending"))).contiguous()
            idx = (torch.arange(b.n, device=be.device) + base).to(torch.int32)
            runs.append((bits, idx))
            base += b.n
        _, order = K.kway_merge(runs)
        return [ColumnBatch.concat(batches).take(order)]
    return [be._cat(inputs[0])]


def _tee(be, node, part, inputs, env):
    b = be._cat(inputs[0])
    return [b for _ in range(node.n_out_ports)] if node.params.get("replicate") else [b]


IDX = "__idx"


def _start_offset(be, inputs) -> int:
    """Global index of this partition's first record: the side input of an indexed operator, produced
    by the count -> exclusive scan stage (K16; LinqToDryad/DryadLinqQueryGen.cs:1225-1248)."""
    if len(inputs) < 2:
        return 0
    side = be._cat(inputs[1])
    return int(side.cols["value"][0].item()) if side.n else 0


def _bind_index(proj: Dict[str, Any], start: int) -> Dict[str, Any]:
    """Replace the symbolic global index of a traced (x, i) lambda by rowid + partition start."""
    gi = X.rowid() + X.lit(int(start))
    return {k: X.wrap(e).substitute({IDX: gi}) for k, e in proj.items()}


def _offsets(be, node, part, inputs, env):
    """Per-partition counts -> exclusive scan -> one start offset per consumer partition (K16)."""
    src = be._cat(inputs[0])
    n = node.params["nports"]
    if "value" not in src.cols or src.cols["value"].dtype != torch.int64:
        return NotImplemented
    counts = torch.zeros(n, dtype=torch.int64, device=be.device)
    counts[:min(n, src.n)] = src.cols["value"][:n]
    offs, _tot = K.exclusive_scan_i64(counts)
    return [ColumnBatch({"value": offs[i:i + 1]}, 1) for i in range(n)]


def _first_false(be, src: ColumnBatch, pred) -> int:
    """Index of the first row for which the predicate is false (src.n when none): one generated
    kernel computes min(rowid where !pred) — no flag column, no scan."""
    if src.n == 0:
        return 0
    big = (1 << 62)
    out, _ = run_on_gpu({"p": X.where(pred, X.lit(big), X.rowid())}, src.cols, src.n, reduce={"p": "min"})
    v = int(out["p"].item())
    return src.n if v >= big else v


def _partition_op(be, node, part, inputs, env):
    """Take / Skip / TakeWhile / SkipWhile on columns (LinqToDryad/DryadLinqVertex.cs:222-289)."""
    p = node.params
    src = be._cat(inputs[0])
    op = p["op"]
    if op == "take":
        return [src.slice(0, max(0, int(p["n"])))]
    if op == "skip":
        return [src.slice(min(src.n, max(0, int(p["n"]))), src.n)]
    lam = p["fn"]
    if src.is_rows or src.has_strings or not isinstance(lam.ir, X.Expr):
        return NotImplemented
    pred = lam.ir
    if p.get("indexed"):
        pred = _bind_index({"p": pred}, 0)["p"]          # stage 2 runs on ONE partition: index = rowid
    cut = _first_false(be, src, pred)
    return [src.slice(0, cut)] if op == "take_while" else [src.slice(cut, src.n)]


def _reverse(be, node, part, inputs, env):
    src = be._cat(inputs[0])
    if src.n <= 1:
        return [src]
    idx = torch.arange(src.n - 1, -1, -1, dtype=torch.int32, device=be.device)
    return [src.take(idx)]


def _zip(be, node, part, inputs, env):
    """Zip over row-aligned partitions (the executor re-aligns the right side to the left side's
    boundaries, LinqToDryad/DryadLinqQueryGen.cs:2626-2717): one fused kernel over both column sets."""
    ir = node.params["fn"].ir
    proj = ir if isinstance(ir, dict) else ({"value": ir} if isinstance(ir, X.Expr) else None)
    if proj is None:
        return NotImplemented
    a, b = be._cat(inputs[0]), be._cat(inputs[1])
    if a.is_rows or b.is_rows or a.has_strings or b.has_strings:
        return NotImplemented
    n = min(a.n, b.n)
    cols = {("a." + k): v[:n].contiguous() for k, v in a.cols.items()}
    cols.update({("b." + k): v[:n].contiguous() for k, v in b.cols.items()})
    if n == 0:
        dt = {"i": torch.int64, "f": torch.float64, "b": torch.bool}
        return [ColumnBatch({k: torch.empty(0, dtype=dt[X.wrap(e).kind], device=be.device) for k, e in proj.items()}, 0,
                            tuple_like=_tuple_cls(ir))]
    out, _ = run_on_gpu(proj, cols, n)
    return [ColumnBatch(out, n, tuple_like=_tuple_cls(ir))]


def _sliding_window(be, node, part, inputs, env):
    """SlidingWindow (LinqToDryad/DryadLinqQueryGen.cs:3234-3293): inputs[1] is the halo — the w-1
    records preceding this partition, delivered by the neighbour (C6).  Window slot j of output row i
    is row i + j of (halo ++ partition): w shifted VIEWS of each column feed one fused kernel."""
    p = node.params
    w = int(p["window"])
    ir = p["fn"].ir
    proj = ir if isinstance(ir, dict) else ({"value": ir} if isinstance(ir, X.Expr) else None)
    if proj is None or w < 1:
        return NotImplemented
    own = be._cat(inputs[0])
    halo = be._cat(inputs[1]) if len(inputs) > 1 and inputs[1] else None
    if own.is_rows or own.has_strings or (halo is not None and not isinstance(halo, ColumnBatch)):
        return NotImplemented
    data = ColumnBatch.concat([halo, own]) if (halo is not None and halo.n) else own
    n_out = max(0, data.n - w + 1)
    dt = {"i": torch.int64, "f": torch.float64, "b": torch.bool}
    if n_out == 0:
        return [ColumnBatch({k: torch.empty(0, dtype=dt[X.wrap(e).kind], device=be.device) for k, e in proj.items()}, 0,
                            tuple_like=_tuple_cls(ir))]
    cols = {}
    for j in range(w):
        for k, v in data.cols.items():
            cols[f"w{j}.{k}"] = v[j:j + n_out]
    out, _ = run_on_gpu(proj, cols, n_out)
    return [ColumnBatch(out, n_out, tuple_like=_tuple_cls(ir))]


def _contains(be, node, part, inputs, env):
    src = be._cat(inputs[0])
    v = node.params["value"]
    if list(src.cols) != ["value"] or src.has_strings or isinstance(v, (str, bytes)) or not isinstance(v, (int, float, bool)):
        return NotImplemented
    if src.n == 0:
        return [ColumnBatch({"value": torch.zeros(1, dtype=torch.bool, device=be.device)}, 1)]
    out, _ = run_on_gpu({"h": X.where(X.col("value", src.schema["value"]) == X.lit(v), 1, 0)}, src.cols, src.n, reduce={"h": "max"})
    return [ColumnBatch({"value": out["h"].reshape(1) > 0}, 1)]


def _proj_of(lam):
    if isinstance(lam.ir, X.Expr):
        return {"value": lam.ir}
    if isinstance(lam.ir, dict):
        return dict(lam.ir)
    return None


def _select(be, node, part, inputs, env):
    p = node.params
    src = be._cat(inputs[0])
    if src.is_rows:
        return NotImplemented
    if p.get("finalize") is not None:
        dec = p["finalize"]
        kinds = [a.out_kind for a in dec.aggs]
        res = dec.result
        proj = {k: _bind_leaves(e, kinds) for k, e in res.items()} if isinstance(res, dict) else {"value": _bind_leaves(res, kinds)}
        # string columns can only be passed through (the key of a string GroupBy); the rest is computed
        passed = {k: src.cols[e.args[0]] for k, e in proj.items()
                  if e.op == "col" and isinstance(src.cols.get(e.args[0]), StrColumn)}
        calc = {k: e for k, e in proj.items() if k not in passed}
        if any(isinstance(src.cols.get(c), StrColumn) for e in calc.values() for c in e.columns()):
            return NotImplemented
        out = run_on_gpu(calc, _numeric_cols(src), src.n)[0] if (calc and src.n) else \
            {k: torch.empty(0, dtype={"i": torch.int64, "f": torch.float64, "b": torch.bool}[e.kind], device=be.device) for k, e in calc.items()}
        return [ColumnBatch({k: (passed[k] if k in passed else out[k]) for k in proj}, src.n, tuple_like=_tuple_cls(res))]
    if src.has_strings:
        return NotImplemented
    proj = _proj_of(p["fn"])
    if proj is None:
        return NotImplemented
    if p["indexed"]:
        proj = _bind_index(proj, _start_offset(be, inputs))
    if src.n == 0:
        dt = {"i": torch.int64, "f": torch.float64, "b": torch.bool}
        return [ColumnBatch({k: torch.empty(0, dtype=dt[X.wrap(e).kind], device=be.device) for k, e in proj.items()}, 0,
                            tuple_like=_tuple_cls(p["fn"].ir))]
    out, _ = run_on_gpu(proj, src.cols, src.n)
    return [ColumnBatch(out, src.n, tuple_like=_tuple_cls(p["fn"].ir))]


def _where(be, node, part, inputs, env):
    p = node.params
    src = be._cat(inputs[0])
    if not isinstance(p["fn"].ir, X.Expr) or src.is_rows:
        return NotImplemented
    if src.n == 0:
        return [src]
    pred = p["fn"].ir
    if p["indexed"]:
        pred = _bind_index({"p": pred}, _start_offset(be, inputs))["p"]
    _, flags = run_on_gpu({}, src.cols, src.n, predicate=pred)
    idx, m = K.compact_flags(flags)
    return [src.take(idx)]


def _key_column(src: ColumnBatch, lam) -> Optional[torch.Tensor]:
    """int64 key-bit column for a traced scalar key (ints: value; floats: IEEE bits)."""
    if not isinstance(lam.ir, X.Expr) or lam.ir.kind not in ("i", "f", "b"):
        return None
    ir = lam.ir
    if ir.op == "col" and ir.args[0] in src.cols:
        t = src.cols[ir.args[0]]
    else:
        out, _ = run_on_gpu({"k": ir}, src.cols, src.n)
        t = out["k"]
    if t.dtype == torch.float64:
        return _canonical_zero(t.view(torch.int64))          # -0.0 == 0.0 as a key
    return t.to(torch.int64) if t.dtype != torch.int64 else t


def _typed_key(src: ColumnBatch, lam, bits: torch.Tensor) -> torch.Tensor:
    """Key column in its numeric dtype (float keys travel as bit patterns in _key_column)."""
    return bits.view(torch.float64) if lam.ir.kind == "f" else bits


def _group_by_sorted(be, node, src: ColumnBatch):
    """Non-decomposable GroupBy (K7): stable radix sort of the key column + run boundaries give the
    groups as CSR on the GPU; one gather brings the rows into group order and the host only slices
    the record list (the reference's ParallelSortGroupBy, LinqToDryad/DryadLinqVertex.cs:6418)."""
    from ..oracle.linq import Grouping
    p = node.params
    if src.is_rows or src.n == 0:
        return NotImplemented
    kb = _key_column(src, p["key"])
    if kb is None:
        return NotImplemented
    order, starts, _keys = K.group_by_sort(kb.contiguous())
    recs = src.take(order).to_records()
    bounds = starts.tolist() + [src.n]
    key_fn = p["key"].fn
    elem = p["elem"].fn if p.get("elem") is not None else None
    res = p["result"].fn if p.get("result") is not None else None
    out = []
    for a, b in zip(bounds, bounds[1:]):
        g = Grouping(key_fn(recs[a]))
        g.extend(recs[a:b] if elem is None else [elem(x) for x in recs[a:b]])
        out.append(res(g.key, g) if res is not None else g)
    return [out]


def _group_by(be, node, part, inputs, env):
    p = node.params
    dec = p.get("decomp")
    if p.get("mode") == "full":
        return _group_by_sorted(be, node, be._cat(inputs[0]))
    if dec is None or p.get("elem") is not None:
        return NotImplemented
    lanes = dec.gpu_lanes()
    if lanes is None:
        return NotImplemented
    src = be._cat(inputs[0])
    mode = p["mode"]
    n = src.n
    skey = src.cols.get("key") if mode == "final_acc" else _str_col(src, p["key"])
    if isinstance(skey, StrColumn):
        return _group_by_str(be, node, src, skey, lanes)
    if src.has_strings:
        return NotImplemented
    pre = _group_by_inputs(be, node, src)
    if pre is NotImplemented:
        return NotImplemented
    kbits, in_views, key_is_f, combine = pre
    cap = max(1024, min(2 * n + 64, be.context.group_capacity_hint))
    # dense integer key range (ids, codes): direct-addressed table, one reduction request per row
    if n >= (1 << 16) and not combine and not key_is_f and len(lanes) < 8:
        lo, hi = K.minmax_i64(kbits.contiguous()).tolist()
        span = hi - lo + 1
        if span <= 4 * min(n, be.context.group_capacity_hint) and span * 64 <= (1 << 30):
            try:
                dt = K.DirectAggTable(lo, span, lanes, be.device)
                dt.update(kbits.contiguous(), [None if c is None else c.contiguous() for c in in_views])
                gk, gl = dt.extract()
                cols_d: Dict[str, torch.Tensor] = {"key": gk}
                for i, a in enumerate(dec.aggs):
                    lane = gl[i].contiguous()
                    cols_d[f"agg{i}"] = lane.view(torch.float64) if lanes[i].endswith("f64") else lane
                return [ColumnBatch(cols_d, gk.numel(), acc_lanes=len(dec.aggs))]
            except OverflowError as ex:
                raise DryadLinqException(ErrorCode.FailureInHashGroupBy, "Arithmetic operation resulted in an overflow.") from ex
    if _has_reserved_key(kbits):
        return NotImplemented
    while True:
        tbl = K.HashAggTable(cap, lanes, be.device)
        if n:
            tbl.update(kbits.contiguous(), [None if c is None else c.contiguous() for c in in_views], combine=combine)
        try:
            gk, gl = tbl.extract()
            break
        except K.N.NativeError as ex:
            if "full" in str(ex) and cap < 4 * n + 1024:
                cap *= 4
                continue
            raise
        except OverflowError as ex:
            raise DryadLinqException(ErrorCode.FailureInHashGroupBy, "Arithmetic operation resulted in an overflow.") from ex
    return [_group_result_batch(dec, lanes, gk, gl, key_is_f)]


def _group_by_str(be, node, src: ColumnBatch, sc: StrColumn, lanes):
    """Decomposable GroupBy on a STRING key: the table is keyed by the 64-bit hash of the bytes, an
    extra Min(row index) lane elects a representative string per group, and every row is then
    compared byte for byte with its group's representative (`tok_verify`), so two different strings
    that share a hash are detected (the vertex then runs on the host path) instead of merged.
    Reference: ParallelHashGroupByPartialAccumulate over string keys with the default comparer
    (LinqToDryad/DryadLinqVertex.cs:5718-6033)."""
    p = node.params
    dec = p["decomp"]
    n = src.n
    combine = p["mode"] == "final_acc"
    if len(lanes) >= 8 or sc.device.type != "cuda":
        return NotImplemented
    if n == 0:
        z = lambda dt: torch.empty(0, dtype=dt, device=be.device)   # noqa: E731
        cols0: Dict[str, Any] = {"key": StrColumn(sc.buf, z(torch.int64), z(torch.int32), z(torch.int64))}
        for i, ln in enumerate(lanes):
            cols0[f"agg{i}"] = z(torch.float64 if ln.endswith("f64") else torch.int64)
        return [ColumnBatch(cols0, 0, acc_lanes=len(dec.aggs))]
    hashes = sc.key_hashes().contiguous()
    if _has_reserved_key(hashes):
        return NotImplemented
    if combine:
        in_cols = [src.cols[f"agg{i}"] for i in range(len(lanes))]
    else:
        proj = {}
        for i, a in enumerate(dec.aggs):
            if a.kind != "count":
                sel = a.selector.ir if a.selector is not None else None
                if not isinstance(sel, X.Expr) or any(isinstance(src.cols.get(c), StrColumn) for c in sel.columns()):
                    return NotImplemented
                proj[f"__v{i}"] = sel
        out = run_on_gpu(proj, _numeric_cols(src), n)[0] if proj else {}
        in_cols = [None if a.kind == "count" else out[f"__v{i}"] for i, a in enumerate(dec.aggs)]
    in_views = [None if c is None else (c.view(torch.int64) if c.dtype == torch.float64 else c.to(torch.int64)).contiguous() for c in in_cols]
    row_idx = torch.arange(n, dtype=torch.int64, device=be.device)
    cap = max(1024, min(2 * n + 64, be.context.group_capacity_hint))
    while True:
        tbl = K.HashAggTable(cap, list(lanes) + ["min_i64"], be.device)
        tbl.update(hashes, in_views + [row_idx], combine=combine)
        try:
            gk, gl = tbl.extract()
            break
        except K.N.NativeError as ex:
            if "full" in str(ex) and cap < 4 * n + 1024:
                cap *= 4
                continue
            raise
        except OverflowError as ex:
            raise DryadLinqException(ErrorCode.FailureInHashGroupBy, "Arithmetic operation resulted in an overflow.") from ex
    rep = gl[len(lanes)]
    rep_start, rep_len = sc.starts[rep].contiguous(), sc.lens[rep].contiguous()
    # full-key verification: row i must equal the representative of the group its hash fell into
    jt = K.JoinTable(gk.contiguous())
    grp = torch.empty(n, dtype=torch.int32, device=be.device)
    matched = torch.empty(n, dtype=torch.uint8, device=be.device)
    K.N.check(K.N.kernels().dryad_join_probe_unique(K.ptr(jt.keys), K.ptr(jt.head), jt.cap, K.ptr(hashes), n,
                                                    K.ptr(grp), K.ptr(matched), K.stream()), "join_probe_unique")
    K._count()
    collisions = K.tokenize_verify(sc.buf, sc.starts.contiguous(), sc.lens.contiguous(),
                                   K.gather_rows(rep_start, grp), K.gather_rows(rep_len, grp))
    if collisions:
        return NotImplemented
    cols: Dict[str, Any] = {"key": StrColumn(sc.buf, rep_start, rep_len, gk)}
    for i in range(len(dec.aggs)):
        lane = gl[i].contiguous()
        cols[f"agg{i}"] = lane.view(torch.float64) if lanes[i].endswith("f64") else lane
    return [ColumnBatch(cols, gk.numel(), acc_lanes=len(dec.aggs))]


def _group_result_batch(dec, lanes, gk, gl, key_is_f) -> ColumnBatch:
    """(group keys, accumulator lanes) -> the partial-aggregate ColumnBatch layout (key, agg0, ...)."""
    cols: Dict[str, torch.Tensor] = {"key": gk.view(torch.float64) if key_is_f else gk}
    for i, a in enumerate(dec.aggs):
        lane = gl[i].contiguous()
        cols[f"agg{i}"] = lane.view(torch.float64) if lanes[i].endswith("f64") else lane
    return ColumnBatch(cols, gk.numel(), acc_lanes=len(dec.aggs))


def _group_by_inputs(be, node, src: ColumnBatch):
    """Key bits and one int64-bits column per accumulator lane for a decomposable GroupBy vertex
    (`mode` partial_acc / full_acc: from the source rows through the traced key and selector
    expressions; final_acc: the (key, agg...) columns of partials).  NotImplemented when a lambda
    did not trace."""
    p = node.params
    dec = p["decomp"]
    lanes = dec.gpu_lanes()
    n = src.n
    if p["mode"] == "final_acc":
        if "key" not in src.cols:
            return NotImplemented
        keys = src.cols["key"]
        key_is_f = keys.dtype == torch.float64
        kbits = keys.view(torch.int64) if key_is_f else keys
        in_cols = [src.cols[f"agg{i}"] for i in range(len(lanes))]
        combine = True
    else:
        if not isinstance(p["key"].ir, X.Expr):
            return NotImplemented
        key_is_f = p["key"].ir.kind == "f"
        proj = {"__k": p["key"].ir}
        for i, a in enumerate(dec.aggs):
            if a.kind != "count":
                sel = a.selector.ir if a.selector is not None else None
                if not isinstance(sel, X.Expr):
                    return NotImplemented
                proj[f"__v{i}"] = sel
        if n:
            out, _ = run_on_gpu(proj, src.cols, n)
        else:
            out = {k: torch.empty(0, dtype=torch.int64, device=be.device) for k in proj}
        kbits = out["__k"].view(torch.int64) if out["__k"].dtype == torch.float64 else out["__k"].to(torch.int64)
        if key_is_f:
            kbits = _canonical_zero(kbits)
        in_cols = [None if a.kind == "count" else out[f"__v{i}"] for i, a in enumerate(dec.aggs)]
        combine = False
    in_views = [None if c is None else (c.view(torch.int64) if c.dtype == torch.float64 else c.to(torch.int64)) for c in in_cols]
    return kbits, in_views, key_is_f, combine


_NEG_ZERO_BITS = -0x8000000000000000


def _canonical_zero(bits: torch.Tensor) -> torch.Tensor:
    """Float keys are compared by bit pattern on the GPU; -0.0 and +0.0 are EQUAL keys, so the sign
    of zero is cleared before the bits are used as a key."""
    return torch.where(bits == _NEG_ZERO_BITS, torch.zeros_like(bits), bits)


def _arena_for(be: GpuBackend, nbytes: int):
    from ..parallel.symm import SymmetricArena
    return SymmetricArena(nbytes + (1 << 20), 0, 1, None, be.device)


def _partition_generic(be, node, src: ColumnBatch, kbits: torch.Tensor, mode: str, seps: Optional[torch.Tensor]):
    """Single-process partitioning with the fused shuffle kernels: all destination partitions
    live on this GPU.  Exact sizes come from a count-only launch, then the scatter runs."""
    from ..parallel.shuffle import ColSpec, FusedShuffle
    nports = node.params["nports"]
    spmd = getattr(be, "spmd_exchange", False) and nports == be.context.world and be.context.world > 1
    # a string column travels as three fixed-width physical columns: key hash, length, and the
    # bytes in zero-padded slots as wide as the longest string (lengths first, bytes behind them)
    phys: Dict[str, torch.Tensor] = {}
    strs = {}
    for k, v in src.cols.items():
        if isinstance(v, StrColumn):
            mx = int(v.lens.max().item()) if src.n else 0
            if spmd:
                import torch.distributed as dist
                sync = be.context.__dict__.get("_spmd_sync")
                if sync is not None:
                    sync()
                t = torch.tensor([mx], dtype=torch.int64, device=be.device)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)      # every rank must use the same slot width
                mx = int(t.item())
            w = max(4, (mx + 3) // 4 * 4)
            if w > 512:
                return None
            strs[k] = w
            phys[k + "#h"] = v.key_hashes()
            phys[k + "#l"] = v.lens.to(torch.int64)
            phys[k + "#b"] = v.pack(w)
        else:
            phys[k] = v
    names = [k for k in phys]
    cols = [kbits.contiguous()] + [phys[k].contiguous() for k in names]
    specs = [ColSpec(8, torch.int64, 1)]
    for k in names:
        t = phys[k]
        if t.dtype == torch.bool:
            t = t.to(torch.uint8)
        w = t.element_size() * (t.shape[1] if t.dim() > 1 else 1)
        if w % 4:
            return None
        specs.append(ColSpec(w, t.dtype, t.shape[1] if t.dim() > 1 else 1))

    def rebuild(d: Dict[str, torch.Tensor]) -> Dict[str, Any]:
        out: Dict[str, Any] = {}
        for k in src.cols:
            out[k] = StrColumn.from_packed(d[k + "#b"], d[k + "#l"], d[k + "#h"]) if k in strs else d[k]
        return out
    n = src.n
    if spmd:
        return _partition_exchange(be, src, names, cols, specs, mode, seps, rebuild)
    cap = n + 16                      # worst case: everything lands in one partition
    total = sum(s.width for s in specs) * cap * nports
    if total > (8 << 30):
        # too skew-tolerant to afford: size for 4x the mean and let the kernel report overflow
        cap = max(1024, 4 * n // nports + 1024)
        total = sum(s.width for s in specs) * cap * nports
    arena = _arena_for(be, total + 4096 * (len(specs) + 2) * nports)
    sh = FusedShuffle(arena, specs, cap, parts_per_rank=nports, emit_key=False, max_rows_in=max(n, 1))
    sh.run(cols, mode=mode, key_col=0, key_kind=0, separators=seps)
    counts = sh.received_rows()
    ports = []
    for j in range(nports):
        m = counts[j]
        d = rebuild({k: sh.recv[c + 1][j][:m] for c, k in enumerate(names)})
        b = ColumnBatch(d, m, src.tuple_like, src.key_spec, src.acc_lanes)
        b._keepalive = (arena, sh)
        ports.append(b)
    return ports


def _partition_exchange(be, src: ColumnBatch, names, cols, specs, mode: str, seps, rebuild=None):
    """SPMD form of a distributor stage (one process per GPU, partition p lives on rank p): the
    partition function, the all-to-all and the consumer's Merge are ONE fused kernel chain writing
    straight into the owners' receive buffers over NVLink (parallel/shuffle.py).  The vertex's
    single output port already holds the rows this rank owns; it is tagged so the executor does not
    exchange it again.  Sizes travel on the control plane (two tiny all-reduces)."""
    import torch.distributed as dist
    from ..parallel.shuffle import FusedShuffle
    ctx = be.context
    n = src.n
    sync = ctx.__dict__.get("_spmd_sync")
    if sync is not None:
        sync()                                           # agreement point before the collective (spmd_control)
    t = torch.tensor([n, -n], dtype=torch.int64, device=be.device)
    tot = torch.tensor([n], dtype=torch.int64, device=be.device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(tot)
    n_max, total = int(t[0].item()), int(tot.item())
    row_bytes = sum(sp.width for sp in specs)
    cap = total + 16                                   # skew-proof when it is affordable
    if cap * row_bytes > (2 << 30):
        cap = max(4096, 2 * total // ctx.world + 4096)   # else 2x the mean; overflow is reported, not silent
    # one FusedShuffle per column layout lives as long as the context (receive buffers, pointer
    # tables and scratch are built once, not per vertex); it is replaced by a larger one on demand
    cache = ctx.__dict__.setdefault("_exchange_cache", {})
    sig = (tuple((sp.width, str(sp.dtype), sp.inner) for sp in specs), mode.startswith("range"))
    sh = cache.get(sig)
    if sh is None or sh.cap_rows < cap or sh._max_in < n_max:
        arena = ctx._spmd_arena()
        grow = 1 if sh is None else 2
        sh = FusedShuffle(arena, specs, cap * grow, parts_per_rank=1, emit_key=False, max_rows_in=max(n_max * grow, 1))
        sh._max_in = max(n_max * grow, 1)
        cache[sig] = sh
    sh.run(cols, mode=mode, key_col=0, key_kind=0, separators=seps)
    try:
        m = sh.received_rows()[0]
    except K.N.NativeError as ex:
        if "overflow" not in str(ex):
            raise
        # skew: the published count matrix (identical on every rank) says how much the fullest owner
        # receives; every rank re-creates the receive side with that room and runs the exchange again
        need = int(sh.counts.sum(dim=0).max().item())
        sh = FusedShuffle(ctx._spmd_arena(), specs, need + 16, parts_per_rank=1, emit_key=False, max_rows_in=sh._max_in)
        sh._max_in = cache[sig]._max_in
        cache[sig] = sh
        sh.run(cols, mode=mode, key_col=0, key_kind=0, separators=seps)
        m = sh.received_rows()[0]
        be.skew_regrows = getattr(be, "skew_regrows", 0) + 1
    # the receive buffers are reused by the next exchange of this layout: hand out copies
    d = {k: sh.recv[c + 1][0][:m].clone() for c, k in enumerate(names)}
    for k in names:
        if k in src.cols and src.cols[k].dtype == torch.bool:
            d[k] = d[k].view(torch.bool) if d[k].dtype == torch.uint8 else d[k]
    if rebuild is not None:
        d = rebuild(d)
    b = ColumnBatch(d, m, src.tuple_like, src.key_spec, src.acc_lanes)
    b._exchanged = True
    be.fused_exchanges = getattr(be, "fused_exchanges", 0) + 1
    return [b]


def _hash_partition(be, node, part, inputs, env):
    src = be._cat(inputs[0])
    if src.is_rows or node.params.get("comparer") is not None:
        return NotImplemented
    sk = _str_col(src, node.params["key"])
    kb = sk.key_hashes() if sk is not None else (None if src.has_strings and not isinstance(node.params["key"].ir, X.Expr)
                                                 else _key_column(src, node.params["key"]))
    if kb is None:
        return NotImplemented
    r = _partition_generic(be, node, src, kb, "hash", None)
    return r if r is not None else NotImplemented


def _sort_bits(t: torch.Tensor) -> torch.Tensor:
    return K.sort_key_bits(t)


def _range_partition(be, node, part, inputs, env):
    p = node.params
    src = be._cat(inputs[0])
    if src.is_rows or p.get("comparer") is not None or not isinstance(p["key"].ir, X.Expr):
        return NotImplemented
    out, _ = run_on_gpu({"k": p["key"].ir}, src.cols, src.n) if src.n else ({"k": torch.empty(0, dtype=torch.int64, device=be.device)}, None)
    kb = _sort_bits(out["k"])
    if p.get("range_keys") is not None:
        sv = torch.tensor(list(p["range_keys"]), dtype=out["k"].dtype if out["k"].numel() else torch.int64, device=be.device)
    else:
        s = be._cat(inputs[1])
        sv = s.cols["value"]
    nports = p["nports"]
    seps = _sort_bits(sv)
    if seps.numel() < nports - 1:      # fewer keys than ports is allowed: trailing ports stay empty
        pad = torch.full((nports - 1 - seps.numel(),), -1 if not p["descending"] else 0, dtype=torch.int64, device=be.device)
        seps = torch.cat([seps, pad])
    mode = "range_desc" if p["descending"] else "range_asc"
    r = _partition_generic(be, node, src, kb, mode, seps.contiguous())
    return r if r is not None else NotImplemented


def _sampler(be, node, part, inputs, env):
    src = be._cat(inputs[0])
    lam = node.params["key"]
    if src.is_rows or not isinstance(lam.ir, X.Expr):
        return NotImplemented
    n = src.n
    if n == 0:
        return [ColumnBatch({"value": torch.empty(0, dtype=torch.int64, device=be.device)}, 0)]
    out, _ = run_on_gpu({"k": lam.ir}, src.cols, n)
    k = out["k"]
    rate = be.context.sample_rate
    cap = int(n * rate * 3) + 1024
    kb = k.view(torch.int64) if k.dtype == torch.float64 else k
    samples, cnt = K.sample_keys(kb.contiguous(), seed=env.vertex_id * 1000 + part, rate=rate, cap=cap)
    m = min(int(cnt.item()), cap)
    if m < 10:                         # too few samples: emit every key (DryadLinqSampler.cs:85-98)
        return [ColumnBatch({"value": k}, n)]
    s = samples[:m]
    return [ColumnBatch({"value": s.view(torch.float64) if k.dtype == torch.float64 else s}, m)]


def _range_sampler(be, node, part, inputs, env):
    p = node.params
    if p.get("comparer") is not None:
        return NotImplemented
    s = be._cat(inputs[0])
    v = s.cols["value"]
    bits = _sort_bits(v).contiguous()
    skeys, _ = K.radix_sort_pairs(bits, None)
    seps_bits = K.select_separators(skeys, skeys.numel(), p["pcount"], bool(p["descending"]))
    if skeys.numel() == 0:
        seps_bits = seps_bits[:0]
    # undo the order-preserving map
    if v.dtype == torch.float64:
        neg = seps_bits >= 0        # original negatives were bit-flipped (top bit clear after map)
        raw = torch.where(neg, ~seps_bits, seps_bits ^ torch.tensor(-0x8000000000000000, dtype=torch.int64, device=be.device))
        out = raw.view(torch.float64)
    else:
        out = seps_bits ^ torch.tensor(-0x8000000000000000, dtype=torch.int64, device=be.device)
    return [ColumnBatch({"value": out}, out.numel())]


_row_sorters: Dict[tuple, Any] = {}


def _order_by(be, node, part, inputs, env):
    p = node.params
    src = be._cat(inputs[0])
    if p.get("comparer") is not None:
        return NotImplemented
    if src.is_rows:
        from ..api.rows import RowKey
        rk = p["key"].fn if isinstance(getattr(p["key"], "fn", None), RowKey) else None
        if rk is None:
            return NotImplemented
        from ..parallel.dsort import RowSorter
        rows = src.cols[ROW_COL]
        sig = (rows.shape[0], rows.shape[1], rk.offset, rk.length, bool(p["descending"]))
        sorter = _row_sorters.get(sig)
        if sorter is None:
            _row_sorters.clear()
            sorter = RowSorter(_arena_for(be, 1 << 20), rows.shape[0], rows.shape[1], key_off=rk.offset,
                               key_len=max(8, rk.length), descending=bool(p["descending"]))
            _row_sorters[sig] = sorter
        out, m = sorter.sort(rows.contiguous())
        return [ColumnBatch.rows(out[:m], rk.offset, rk.length)]
    if not isinstance(p["key"].ir, X.Expr):
        return NotImplemented
    if src.n <= 1:
        return [src]
    out, _ = run_on_gpu({"k": p["key"].ir}, src.cols, src.n)
    bits = K.sort_key_bits(out["k"], descending=bool(p["descending"])).contiguous()
    _, idx = K.radix_sort_pairs(bits, None, iota=True)
    return [src.take(idx)]


def _checked_sum_i64(col: torch.Tensor) -> int:
    """Exact Sum of an int64 column with C# checked-arithmetic semantics (LinqToDryad/
    DryadLinqDecomposition.cs:312-315): the low and high 32-bit halves are summed separately in ONE
    generated kernel (neither partial sum can wrap for n < 2^31), recombined as an exact integer on
    the host, and an OverflowException is raised when the result leaves the int64 range."""
    n = col.numel()
    if n == 0:
        return 0
    if n >= (1 << 31):
        raise CompileError("column too long for the split checked sum")
    c = X.col("v", "i")
    out, _ = run_on_gpu({"lo": c & X.lit(0xFFFFFFFF), "hi": c >> X.lit(32)}, {"v": col.contiguous()}, n,
                        reduce={"lo": "sum", "hi": "sum"})
    total = (int(out["hi"].item()) << 32) + int(out["lo"].item())
    if not -(1 << 63) <= total < (1 << 63):
        raise DryadLinqException(ErrorCode.FailureInOperator, "Arithmetic operation resulted in an overflow.")
    return total


def _scalar(be, v, dtype) -> ColumnBatch:
    return ColumnBatch({"value": torch.tensor([v], dtype=dtype, device=be.device)}, 1)


def _basic_aggregate(be, node, part, inputs, env):
    """Scalar aggregates on columns (K11; LinqToDryad/DryadLinqVertex.cs:1673-4704): stage 1 reduces
    a partition to one partial, stage 2 combines the partials of all partitions."""
    p = node.params
    kind, stage = p["kind"], p["stage"]
    if kind not in ("count", "sum", "min", "max", "average", "any", "all", "first", "last", "any_true"):
        return NotImplemented
    src = be._cat(inputs[0])
    if src.is_rows:
        if kind == "count" and stage == 1 and p.get("fn") is None:
            return [_scalar(be, src.n, torch.int64)]
        return NotImplemented
    if src.has_strings:
        return NotImplemented
    fn = p.get("fn")
    if fn is not None and not isinstance(fn.ir, X.Expr):
        return NotImplemented
    if kind == "any_true":                                  # stage 2 of Contains
        if "value" not in src.cols:
            return NotImplemented
        return [_scalar(be, bool(src.n and bool(src.cols["value"].any().item())), torch.bool)]
    if stage == 1:
        n = src.n
        if kind in ("count", "any", "all"):
            if fn is None:
                hits = n
            elif n == 0:
                hits = 0
            else:
                _, cnt = run_on_gpu({}, src.cols, n, predicate=fn.ir, count_flags=True)
                hits = int(cnt.item())
            if kind == "count":
                return [_scalar(be, hits, torch.int64)]
            return [_scalar(be, hits > 0 if kind == "any" else hits == n, torch.bool)]
        if kind in ("first", "last"):
            if n == 0:
                return [src]
            if fn is None:
                return [src.slice(0, 1) if kind == "first" else src.slice(n - 1, n)]
            big = 1 << 62
            if kind == "first":
                out, _ = run_on_gpu({"p": X.where(fn.ir, X.rowid(), X.lit(big))}, src.cols, n, reduce={"p": "min"})
                i = int(out["p"].item())
                return [src.slice(i, i + 1) if i < big else src.slice(0, 0)]
            out, _ = run_on_gpu({"p": X.where(fn.ir, X.rowid(), X.lit(-1))}, src.cols, n, reduce={"p": "max"})
            i = int(out["p"].item())
            return [src.slice(i, i + 1) if i >= 0 else src.slice(0, 0)]
        # sum / min / max / average over a value column or a traced selector
        if fn is not None:
            empty_dt = torch.float64 if getattr(fn.ir, "kind", "i") == "f" else torch.int64
            out, _ = run_on_gpu({"v": fn.ir}, src.cols, n) if n else ({"v": torch.empty(0, dtype=empty_dt, device=be.device)}, None)
            col = out["v"]
        else:
            if "value" not in src.cols:
                return NotImplemented
            col = src.cols["value"]
        if col.dtype == torch.bool:
            return NotImplemented
        if n == 0:
            if kind == "sum":
                return [ColumnBatch({"value": torch.zeros(1, dtype=col.dtype, device=be.device)}, 1)]
            if kind == "average":                          # an empty partition's partial has the (sum, count) shape too
                return [ColumnBatch({"item1": col[:0], "item2": torch.empty(0, dtype=torch.int64, device=be.device)}, 0, tuple_like=True)]
            return [ColumnBatch({"value": col[:0]}, 0)]
        if kind in ("sum", "average"):
            tot = _checked_sum_i64(col) if col.dtype == torch.int64 else K.reduce(col.contiguous(), "sum")
            if kind == "sum":
                return [_scalar(be, tot, torch.int64) if col.dtype == torch.int64 else ColumnBatch({"value": tot.reshape(1)}, 1)]
            pair = ColumnBatch({"item1": torch.tensor([tot], dtype=torch.int64, device=be.device) if col.dtype == torch.int64 else tot.reshape(1),
                                "item2": torch.tensor([n], dtype=torch.int64, device=be.device)}, 1, tuple_like=True)
            return [pair]
        return [ColumnBatch({"value": K.reduce(col.contiguous(), kind).reshape(1)}, 1)]
    # ---- stage 2 over the per-partition partials --------------------------------------------------
    if kind in ("first", "last"):
        if src.n == 0:
            return NotImplemented                              # default / typed "no elements" error: host
        return [src.slice(0, 1) if kind == "first" else src.slice(src.n - 1, src.n)]
    if kind == "average":
        if src.n == 0 or "item1" not in src.cols:
            return NotImplemented
        s_col, c_col = src.cols["item1"], src.cols["item2"]
        cnt = int(K.reduce(c_col.contiguous(), "sum").item())
        if cnt == 0:
            return NotImplemented
        tot = _checked_sum_i64(s_col) if s_col.dtype == torch.int64 else float(K.reduce(s_col.contiguous(), "sum").item())
        return [_scalar(be, tot / cnt, torch.float64)]
    if "value" not in src.cols:
        return NotImplemented
    col = src.cols["value"]
    if kind in ("any", "all"):
        if src.n == 0:
            return [_scalar(be, kind == "all", torch.bool)]
        v = bool(col.any().item()) if kind == "any" else bool(col.all().item())
        return [_scalar(be, v, torch.bool)]
    if src.n == 0:
        if kind in ("count", "sum"):
            return [_scalar(be, 0, torch.int64)]
        return NotImplemented          # host raises the typed "no elements" error
    if kind in ("count", "sum") and col.dtype == torch.int64:
        return [_scalar(be, _checked_sum_i64(col), torch.int64)]
    op = "sum" if kind in ("count", "sum") else kind
    return [ColumnBatch({"value": K.reduce(col.contiguous(), op).reshape(1)}, 1)]


def _join(be, node, part, inputs, env):
    p = node.params
    if node.type != NodeType.Join or p.get("comparer") is not None:
        return NotImplemented
    use_merge = p.get("algorithm") == "merge"
    res = p["result"].ir
    if not isinstance(res, (dict, X.Expr)):
        return NotImplemented
    outer, inner = be._cat(inputs[0]), be._cat(inputs[1])
    if outer.is_rows or inner.is_rows:
        return NotImplemented
    ok, ik = _key_column(outer, p["outer_key"]), _key_column(inner, p["inner_key"])
    if ok is None or ik is None:
        return NotImplemented
    if p["outer_key"].ir.kind != p["inner_key"].ir.kind:
        # keys are matched by bit pattern: a float key never equals an int key's bits, while the
        # host path compares values (1.0 == 1); mixed kinds run there
        return NotImplemented
    if outer.n == 0 or inner.n == 0:
        proj = res if isinstance(res, dict) else {"value": res}
        dt = {"i": torch.int64, "f": torch.float64, "b": torch.bool}
        return [ColumnBatch({k: torch.empty(0, dtype=dt[e.kind], device=be.device) for k, e in proj.items()}, 0,
                            tuple_like=_tuple_cls(res))]
    if use_merge:                                        # both inputs are key-sorted: merge join (K9)
        desc = bool(p.get("descending"))
        okb = K.sort_key_bits(_typed_key(outer, p["outer_key"], ok), descending=desc).contiguous()
        ikb = K.sort_key_bits(_typed_key(inner, p["inner_key"], ik), descending=desc).contiguous()
        op, ib = K.merge_join(okb, ikb)
    else:
        if _has_reserved_key(ik, ok):
            return NotImplemented
        jt = K.JoinTable(ik.contiguous())                # build on the inner (usually smaller) side
        fused = _join_fused(be, jt, outer, inner, ok, res)
        if fused is not None:
            return fused
        op, ib = jt.probe(ok.contiguous())
    o_sel, i_sel = outer.take(op), inner.take(ib)
    cols = {("o." + k): v for k, v in o_sel.cols.items()}
    cols.update({("i." + k): v for k, v in i_sel.cols.items()})
    proj = res if isinstance(res, dict) else {"value": res}
    out, _ = run_on_gpu(proj, cols, op.numel()) if op.numel() else ({}, None)
    if not op.numel():
        dt = {"i": torch.int64, "f": torch.float64, "b": torch.bool}
        out = {k: torch.empty(0, dtype=dt[e.kind], device=be.device) for k, e in proj.items()}
    return [ColumnBatch(out, op.numel(), tuple_like=_tuple_cls(res))]


def _group_join(be, node, part, inputs, env):
    """GroupJoin whose result selector uses the group through decomposable aggregates
    (ParallelHashGroupJoin, LinqToDryad/DryadLinqVertex.cs:7316-7761): the inner side is aggregated
    by key in the hash-aggregate table, every outer row probes the group table once, and the selector
    runs as one fused kernel over the outer columns and the gathered aggregates.  Outer rows without
    a match see the empty group (Count 0, Sum 0)."""
    p = node.params
    dec = p.get("gj_decomp")
    if dec is None or p.get("comparer") is not None or p.get("algorithm") == "merge":
        return NotImplemented
    lanes = dec.gpu_lanes()
    if lanes is None or len(lanes) >= 8:
        return NotImplemented
    outer, inner = be._cat(inputs[0]), be._cat(inputs[1])
    if outer.is_rows or inner.is_rows or outer.has_strings or inner.has_strings:
        return NotImplemented
    if not isinstance(p["outer_key"].ir, X.Expr) or not isinstance(p["inner_key"].ir, X.Expr) or \
            p["outer_key"].ir.kind != p["inner_key"].ir.kind:
        return NotImplemented
    ok, ik = _key_column(outer, p["outer_key"]), _key_column(inner, p["inner_key"])
    if ok is None or ik is None or _has_reserved_key(ok, ik):
        return NotImplemented
    res = dec.result
    kinds = [a.out_kind for a in dec.aggs]
    proj = {k: _bind_leaves(e, kinds) for k, e in res.items()} if isinstance(res, dict) else {"value": _bind_leaves(res, kinds)}
    dt = {"i": torch.int64, "f": torch.float64, "b": torch.bool}
    if outer.n == 0:
        return [ColumnBatch({k: torch.empty(0, dtype=dt[e.kind], device=be.device) for k, e in proj.items()}, 0,
                            tuple_like=_tuple_cls(res))]
    # 1. aggregate the inner side by key
    sel_proj = {}
    for i, a in enumerate(dec.aggs):
        if a.kind != "count":
            sel = a.selector.ir if a.selector is not None else None
            if not isinstance(sel, X.Expr):
                return NotImplemented
            sel_proj[f"__v{i}"] = sel
    empty_ok = all(a.kind in ("count", "sum") for a in dec.aggs)
    cols = {("o." + k): v for k, v in outer.cols.items()}
    if inner.n:
        vals = run_on_gpu(sel_proj, inner.cols, inner.n)[0] if sel_proj else {}
        in_views = []
        for i, a in enumerate(dec.aggs):
            if a.kind == "count":
                in_views.append(None)
            else:
                c = vals[f"__v{i}"]
                in_views.append((c.view(torch.int64) if c.dtype == torch.float64 else c.to(torch.int64)).contiguous())
        tbl = K.HashAggTable(max(1024, 2 * inner.n + 64), lanes, be.device)
        try:
            tbl.update(ik.contiguous(), in_views)
            gk, gl = tbl.extract()
        except OverflowError as ex:
            raise DryadLinqException(ErrorCode.FailureInHashGroupBy, "Arithmetic operation resulted in an overflow.") from ex
        # 2. one probe per outer row, 3. gather the group's aggregates (identity where unmatched)
        jt = K.JoinTable(gk.contiguous())
        rows = torch.empty(outer.n, dtype=torch.int32, device=be.device)
        matched = torch.empty(outer.n, dtype=torch.uint8, device=be.device)
        K.N.check(K.N.kernels().dryad_join_probe_unique(K.ptr(jt.keys), K.ptr(jt.head), jt.cap, K.ptr(ok.contiguous()), outer.n,
                                                        K.ptr(rows), K.ptr(matched), K.stream()), "join_probe_unique")
        K._count()
        if not empty_ok and not bool(matched.all().item()):
            return NotImplemented                             # Min/Max/Average of an empty group: host semantics
        rows = rows.clamp_(min=0)
        m64 = matched.to(torch.int64)
        for i, ln in enumerate(lanes):
            lane = K.gather_rows(gl[i].contiguous(), rows)
            if ln.endswith("f64"):
                cols[f"agg{i}"] = lane.view(torch.float64) * matched.to(torch.float64)
            else:
                cols[f"agg{i}"] = lane * m64
    else:
        if not empty_ok:
            return NotImplemented
        for i, ln in enumerate(lanes):
            cols[f"agg{i}"] = torch.zeros(outer.n, dtype=torch.float64 if ln.endswith("f64") else torch.int64, device=be.device)
    out, _ = run_on_gpu(proj, cols, outer.n)
    return [ColumnBatch(out, outer.n, tuple_like=_tuple_cls(res))]


def _join_fused(be, jt, outer: ColumnBatch, inner: ColumnBatch, ok: torch.Tensor, res):
    """Unique-key build side (dimension table): probe, inner-column gather and the result selector
    run as ONE generated kernel over the outer rows — no (probe, build) pair lists, no gathered
    intermediates; matched rows are then compacted.  Returns None when not applicable."""
    try:
        if not jt.unique:
            return None
        proj = res if isinstance(res, dict) else {"value": res}
        hit = X.probe_slot("jt", X.col("__okey"))          # slot, not row: payloads are permuted to slot order below
        sub = {}
        arrays = {}
        for k, v in inner.cols.items():
            if v.dim() != 1:
                return None
            kind = "f" if v.dtype in (torch.float32, torch.float64) else ("b" if v.dtype == torch.bool else "i")
            g = X.gather("i." + k, hit, "f" if kind == "f" else "i")
            sub["i." + k] = X.cast(g, "b") if kind == "b" else g
            arrays["i." + k] = (v if v.element_size() in (4, 8) else v.to(torch.int64)).contiguous()   # slot_payload gathers 4/8-byte elements
        outs = {}
        for name, e in proj.items():
            e = X.wrap(e)
            if e.kind not in ("i", "f", "b"):
                return None
            used = {c for c in e.columns() if c.startswith("i.")}
            outs[name] = e.substitute({c: sub[c] for c in used})
        cols = {("o." + k): v for k, v in outer.cols.items()}
        cols["__okey"] = ok.contiguous()
        need = set()
        for e in outs.values():
            need |= {a for a in arrays if _mentions_array(e, a)}
        out, flags = run_on_gpu(outs, cols, outer.n, predicate=hit >= 0, tables={"jt": jt},
                                arrays={a: jt.slot_payload(arrays[a]) for a in need})
    except (CompileError, X.TraceError):
        return None
    idx, m = K.compact_flags(flags)
    batch = ColumnBatch(out, outer.n, tuple_like=_tuple_cls(res))
    return [batch.take(idx)]


def _mentions_array(e, name: str) -> bool:
    if not isinstance(e, X.Expr):
        return False
    if e.op == "gather" and e.args[0] == name:
        return True
    return any(_mentions_array(a, name) for a in e.args)


def _distinct(be, node, part, inputs, env):
    src = be._cat(inputs[0])
    if list(src.cols) != ["value"] or src.cols["value"].dtype not in (torch.int64, torch.float64):
        return NotImplemented
    v = src.cols["value"]
    if src.n == 0:
        return [src]
    kb = v.view(torch.int64) if v.dtype == torch.float64 else v
    if _has_reserved_key(kb):
        return NotImplemented
    tbl = K.HashAggTable(2 * src.n + 64, [], be.device)
    tbl.update(kb.contiguous(), [])
    gk, _ = tbl.extract()
    return [ColumnBatch({"value": gk.view(torch.float64) if v.dtype == torch.float64 else gk}, gk.numel())]


def _set_operation(be, node, part, inputs, env):
    a, b = be._cat(inputs[0]), be._cat(inputs[1])
    ok = lambda x: (not x.is_rows) and list(x.cols) == ["value"] and x.cols["value"].dtype in (torch.int64, torch.float64)   # noqa: E731
    if not ok(a) or not ok(b) or a.cols["value"].dtype != b.cols["value"].dtype:
        return NotImplemented
    f = a.cols["value"].dtype == torch.float64
    ab = a.cols["value"].view(torch.int64) if f else a.cols["value"]
    bb = b.cols["value"].view(torch.int64) if f else b.cols["value"]
    if _has_reserved_key(ab, bb):
        return NotImplemented
    r = K.set_operation(node.params["op"], ab.contiguous(), bb.contiguous())
    return [ColumnBatch({"value": r.view(torch.float64) if f else r}, r.numel())]


# operators that understand string columns (StrColumn); the others see them only through the host path
_STR_AWARE = {NodeType.Apply, NodeType.InputTable, NodeType.OutputTable, NodeType.Dummy, NodeType.Merge, NodeType.Concat, NodeType.Tee,
              NodeType.ForkBranch, NodeType.Select, NodeType.SelectMany, NodeType.GroupBy, NodeType.HashPartition}

_GPU_OPS = {
    NodeType.SelectMany: _select_many, NodeType.Apply: _apply, NodeType.Offsets: _offsets, NodeType.PartitionOp: _partition_op,
    NodeType.Reverse: _reverse, NodeType.Zip: _zip, NodeType.Contains: _contains, NodeType.SlidingWindow: _sliding_window,
    NodeType.InputTable: _input, NodeType.OutputTable: _passthrough, NodeType.Dummy: _passthrough,
    NodeType.Merge: _passthrough, NodeType.Concat: _passthrough, NodeType.ForkBranch: _passthrough,
    NodeType.Tee: _tee, NodeType.Select: _select, NodeType.Where: _where, NodeType.GroupBy: _group_by,
    NodeType.HashPartition: _hash_partition, NodeType.RangePartition: _range_partition,
    NodeType.Sampler: _sampler, NodeType.RangeSampler: _range_sampler, NodeType.OrderBy: _order_by,
    NodeType.BasicAggregate: _basic_aggregate, NodeType.Join: _join, NodeType.GroupJoin: _group_join, NodeType.Distinct: _distinct,
    NodeType.SetOperation: _set_operation,
}
