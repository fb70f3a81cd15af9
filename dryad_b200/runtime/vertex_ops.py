"""Host vertex runtime: the executable semantics of every physical operator over Python record
lists — the role of the reference's operator library DryadLinqVertex
(LinqToDryad/DryadLinqVertex.cs:51-5311) and helper bodies (LinqToDryad/DryadLinqHelper.cs:38-458).

This is the general path (any Python record type, opaque lambdas).  Stages whose lambdas traced
to column expressions over numeric columns run on the GPU through `runtime/gpu_ops.py` instead;
both produce the same records, which is what the differential tests check.

Contract: `run(node, part_index, inputs, env) -> [port0, port1, ...]` where `inputs[k]` is the
list of channel payloads arriving on child edge k (one payload per producer for merging edges).
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Sequence

from ..api.errors import DryadLinqException, ErrorCode
from ..oracle import linq as L
from ..oracle.sampler import phase1_sample, range_bucket, select_separators
from ..plan.qnodes import DLinqQueryNode, NodeType
from ..utils.hashing import hash_bucket


def _cat(payloads: Sequence[Sequence]) -> List:
    out: List = []
    for p in payloads:
        out.extend(p)
    return out


def _kfn(lam) -> Callable:
    return lam.fn if lam is not None and lam.fn is not None else (lambda x: x)


class VertexEnv:
    """Per-vertex environment (LinqToDryad/VertexEnv.cs:34-367): ids, partition index, context,
    thread count, write-buffer sizing and the error report file."""

    DEFAULT_WRITE_BUFFER = 8 << 20          # VertexEnv.cs:261-283
    LARGE_WRITE_BUFFER = 256 << 20
    MIN_WRITE_BUFFER = 16 << 10

    def __init__(self, context, vertex_id: int, part: int, version: int = 0, n_outputs: int = 1,
                 use_large_buffer: bool = False):
        self.context, self.vertex_id, self.part, self.version = context, vertex_id, part, version
        self.n_outputs = max(1, n_outputs)
        from ..plan import static_config as SC
        self.use_large_buffer = use_large_buffer or SC.USE_LARGE_BUFFER
        self._log = None

    @property
    def log(self):
        """`vertex-<id>-<version>-LinqLog.txt` (LinqToDryad/DryadLinqLog.cs:30-190), created on first use."""
        if self._log is None:
            import os
            from ..utils.log import DryadLinqLog
            level = getattr(getattr(self.context, "runtime_logging_level", None), "name", None)
            self._log = DryadLinqLog.for_vertex(os.path.join(getattr(self.context, "work_dir", "."), "log"),
                                                self.vertex_id, self.version, level)
        return self._log

    @property
    def threads_per_worker(self) -> int:
        """DRYAD_THREADS_PER_WORKER, else the context setting, else the core count (VertexEnv.cs:72-86)."""
        import os
        v = os.environ.get("DRYAD_THREADS_PER_WORKER")
        if v and v.isdigit() and int(v) > 0:
            return int(v)
        t = getattr(self.context, "threads_per_worker", 0) or 0
        return t if t > 0 else (os.cpu_count() or 1)

    def write_buffer_size(self, available_physical: int = 0) -> int:
        """8 MB per output (256 MB when a large buffer is requested), capped by a quarter of the
        available physical memory shared among the outputs, never below 16 KB (GetWriteBuffSize)."""
        want = self.LARGE_WRITE_BUFFER if self.use_large_buffer else self.DEFAULT_WRITE_BUFFER
        if available_physical <= 0:
            try:
                import psutil
                available_physical = int(psutil.virtual_memory().available)
            except Exception:  # noqa: BLE001
                available_physical = 0
        if available_physical > 0:
            want = min(want, available_physical // 4 // self.n_outputs)
        return max(self.MIN_WRITE_BUFFER, want)

    def report_vertex_error(self, exc: BaseException, directory: str = None) -> str:
        """Write the failure to `VertexException.txt` in the vertex's working directory so the
        graph manager / job browser can show it (ReportVertexError, VertexEnv.cs:230-259)."""
        import os
        import traceback
        d = directory or getattr(self.context, "work_dir", ".")
        os.makedirs(d, exist_ok=True)
        path = os.path.join(d, "VertexException.txt")
        with open(path, "w", encoding="utf-8") as f:
            f.write(f"vertex {self.vertex_id} partition {self.part} version {self.version}\n")
            f.write("".join(traceback.format_exception(type(exc), exc, exc.__traceback__)))
        self.log.error("vertex {0}[{1}] v{2} failed: {3}", self.vertex_id, self.part, self.version, repr(exc))
        return path


def _native_split_count(members, inputs, env):
    """Traced pipeline [SelectMany(l => l.Split(c)), GroupBy(w => w, partial, Count only)] over raw text
    partitions: the whole vertex body runs in the native host library (csrc/host/vertex_native.cpp
    `hv_split_count`) instead of over Python records.  Returns the partial-aggregate records
    [(word, [count])] and the number of members consumed, or None when the pattern does not apply."""
    from ..expr import ir as X
    if len(members) < 2 or members[0].type != NodeType.SelectMany or members[1].type != NodeType.GroupBy:
        return None
    sm, gb = members[0].params, members[1].params
    ir = sm["fn"].ir
    if sm.get("indexed") or sm.get("result") is not None or not isinstance(ir, X.Expr) or ir.op != "split":
        return None
    dec = gb.get("decomp")
    key_ir = getattr(gb.get("key"), "ir", None)
    if dec is None or gb.get("mode") not in ("partial_acc", "full_acc") or gb.get("elem") is not None or gb.get("comparer") is not None:
        return None
    if not (isinstance(key_ir, X.Expr) and key_ir.op == "col" and key_ir.kind == "s"):
        return None
    if len(dec.aggs) != 1 or dec.aggs[0].kind != "count":
        return None
    parts = [p for p in inputs[0]]
    if not parts or not all(isinstance(getattr(p, "data", None), (bytes, bytearray)) and getattr(p, "device_buf", None) is None
                            for p in parts):
        return None
    if any(b"\r" in p.data for p in parts):
        return None                                   # CRLF text: the record reader strips the '\r'
    import ctypes as C
    from .. import _native as N
    lib = N.host()
    delim = ir.args[1].encode("utf-8")
    merged: dict = {}
    for p in parts:
        h = lib.hv_split_count(p.data, len(p.data), delim[0], 1, env.threads_per_worker)
        try:
            g, nb = lib.hv_groups(h), lib.hv_word_bytes(h)
            words = C.create_string_buffer(max(nb, 1))
            offs = (C.c_int64 * (g + 1))()
            cnts = (C.c_int64 * max(g, 1))()
            lib.hv_export(h, words, offs, cnts)
        finally:
            lib.hv_free(h)
        raw = words.raw
        for i in range(g):
            w = raw[offs[i]:offs[i + 1]].decode("utf-8")
            merged[w] = merged.get(w, 0) + cnts[i]
    return [(w, [c]) for w, c in merged.items()], 2


def run(node: DLinqQueryNode, part: int, inputs: List[List], env: VertexEnv) -> List[List]:
    if node.type == NodeType.Super:
        cur = inputs
        out: List[List] = []
        members = list(node.members)
        fast = _native_split_count(members, inputs, env)
        if fast is not None:
            recs, used = fast
            out = [recs]
            cur = [[recs]]
            members = members[used:]
            env.native_vertices = getattr(env, "native_vertices", 0) + 1
            env.context.__dict__["native_host_vertices"] = env.context.__dict__.get("native_host_vertices", 0) + 1
        for m in members:
            out = run(m, part, cur, env)
            cur = [[out[0]]]
        return out
    fn = _DISPATCH.get(node.type)
    if fn is None:
        raise DryadLinqException(ErrorCode.OperatorNotSupported, f"no vertex code for {node.type}")
    return fn(node, part, inputs, env)


# ---- sources / sinks / plumbing ----------------------------------------------------------------
def _input(node, part, inputs, env):
    return [env.context._read_table_partition(node.params["table"], part)]


def _output(node, part, inputs, env):
    return [_cat(inputs[0])]


def _identity(node, part, inputs, env):
    return [_cat(inputs[0])]


def _merge(node, part, inputs, env):
    p = node.params
    if p.get("sort_key") is not None:
        import heapq
        key = _kfn(p["sort_key"])
        runs = [list(r) for r in inputs[0]]
        return [list(heapq.merge(*runs, key=key, reverse=bool(p.get("descending"))))]
    if p.get("keep_port_order", True):
        return [_cat(inputs[0])]       # port order is preserved (keepPortOrder)
    # unordered merge: ports are read in a permutation seeded by the vertex id (VertexReader.cs:67-83)
    from ..channel.factory import port_order
    ports = list(inputs[0])
    return [_cat([ports[i] for i in port_order(len(ports), False, env.vertex_id * 7919 + part)])]


def _tee(node, part, inputs, env):
    data = _cat(inputs[0])
    if node.params.get("replicate"):
        return [data for _ in range(node.n_out_ports)]
    return [data]


def _concat(node, part, inputs, env):
    return [_cat(inputs[0])]


def _fork_branch(node, part, inputs, env):
    return [_cat(inputs[0])]


# ---- element-wise ---------------------------------------------------------------------------------
def _start_index(inputs) -> int:
    side = _cat(inputs[1])
    return int(side[0]) if side else 0


def _select(node, part, inputs, env):
    p = node.params
    src = _cat(inputs[0])
    if p.get("finalize") is not None:
        dec = p["finalize"]
        return [[dec.final(k, accs) for k, accs in src]]
    if p["indexed"]:
        s = _start_index(inputs)
        f = p["fn"].fn
        return [[f(x, s + i) for i, x in enumerate(src)]]
    return [L.select(src, p["fn"].fn)]


def _where(node, part, inputs, env):
    p = node.params
    src = _cat(inputs[0])
    if p["indexed"]:
        s = _start_index(inputs)
        f = p["fn"].fn
        return [[x for i, x in enumerate(src) if f(x, s + i)]]
    return [L.where(src, p["fn"].fn)]


def _select_many(node, part, inputs, env):
    p = node.params
    src = _cat(inputs[0])
    res = p["result"].fn if p.get("result") else None
    if p["indexed"]:
        s = _start_index(inputs)
        out = []
        for i, x in enumerate(src):
            for y in p["fn"].fn(x, s + i):
                out.append(res(x, y) if res else y)
        return [out]
    return [L.select_many(src, p["fn"].fn, res)]


# ---- distributors -----------------------------------------------------------------------------------
def _hash_partition(node, part, inputs, env):
    p = node.params
    n = p["nports"]
    key = _kfn(p["key"])
    ports: List[List] = [[] for _ in range(n)]
    for x in _cat(inputs[0]):
        ports[hash_bucket(key(x), n)].append(x)       # (hash & 0x7FFFFFFF) % nPorts
    return ports


def _range_partition(node, part, inputs, env):
    p = node.params
    n = p["nports"]
    key = _kfn(p["key"])
    seps = p["range_keys"] if p.get("range_keys") is not None else _cat(inputs[1])
    if len(seps) > n - 1:
        raise DryadLinqException(ErrorCode.RangePartitionInputOutputMismatch,
                                 f"{len(seps)} range keys for {n} ports")
    desc = bool(p["descending"])
    cmp = p.get("comparer")
    ports: List[List] = [[] for _ in range(n)]
    if cmp is None:
        for x in _cat(inputs[0]):
            ports[range_bucket(key(x), seps, desc)].append(x)
    else:
        for x in _cat(inputs[0]):
            k = key(x)
            lo, hi = 0, len(seps)
            while lo < hi:
                mid = (lo + hi) // 2
                c = cmp(seps[mid], k)
                if (c > 0) if desc else (c < 0):
                    lo = mid + 1
                else:
                    hi = mid
            ports[lo].append(x)
    return ports


def _sampler(node, part, inputs, env):
    key = _kfn(node.params["key"])
    return [phase1_sample([key(x) for x in _cat(inputs[0])], seed=env.vertex_id * 1000 + part,
                          rate=env.context.sample_rate)]


def _range_sampler(node, part, inputs, env):
    p = node.params
    samples = _cat(inputs[0])
    cmp = p.get("comparer")
    if cmp is not None:
        import functools
        samples = sorted(samples, key=functools.cmp_to_key(cmp))
    else:
        samples = sorted(samples)
    return [select_separators(samples, p["pcount"], bool(p["descending"]))]


def _offsets(node, part, inputs, env):
    counts = _cat(inputs[0])
    offs = L.offsets(counts)
    n = node.params["nports"]
    return [[offs[i]] if i < len(offs) else [0] for i in range(n)]


# ---- stateful operators --------------------------------------------------------------------------------
def _order_by(node, part, inputs, env):
    p = node.params
    return [L.order_by(_cat(inputs[0]), _kfn(p["key"]), bool(p["descending"]), p.get("comparer"))]


_PLAIN_KEYS = (str, int, float, bool, bytes)


def _is_identity(fn) -> bool:
    probe = object()
    try:
        return fn(probe) is probe
    except Exception:  # noqa: BLE001
        return False


def _fast_accumulate(src, key, elem, dec):
    """Tight loops for the common decomposable shapes (Count / Sum / Min / Max over plain hashable
    keys): the generic loop pays four Python calls per record (key wrapper, _hashable, seed or
    accumulate per aggregate); these pay one or two.  Count with an identity key is a C-speed
    `collections.Counter`.  Returns None when the shape does not qualify (generic loop runs).
    Same results as the generic loop, including checked integer sums (Python ints do not wrap;
    the int64 range check happens where the generic path does it: at combine / final time)."""
    aggs = dec.aggs
    if elem is not None or not src or any(a.kind not in ("count", "sum", "min", "max") for a in aggs):
        return None
    first = src[0]
    if not isinstance(key(first), _PLAIN_KEYS):
        return None
    kinds = [a.kind for a in aggs]
    sels = [a.selector.fn if a.selector is not None else None for a in aggs]
    ident = _is_identity(key)
    if kinds == ["count"]:
        from collections import Counter
        c = Counter(src) if ident else Counter(map(key, src))
        if not all(isinstance(k, _PLAIN_KEYS) for k in c):
            return None
        return [(k, [n]) for k, n in c.items()]
    from ..plan.decomposition import checked_add
    table: Dict[Any, list] = {}
    get = table.get
    na = len(aggs)
    for x in src:
        k = x if ident else key(x)
        if type(k) not in _PLAIN_KEYS:
            return None                                   # mixed key types: let the generic loop decide
        accs = get(k)
        if accs is None:
            table[k] = [1 if kinds[j] == "count" else (sels[j](x) if sels[j] is not None else x) for j in range(na)]
            continue
        for j in range(na):
            kd = kinds[j]
            if kd == "count":
                accs[j] += 1
                continue
            v = sels[j](x) if sels[j] is not None else x
            if kd == "sum":
                a = accs[j]
                accs[j] = checked_add(a, v) if type(a) is int and type(v) is int else a + v
            elif kd == "min":
                if v < accs[j]:
                    accs[j] = v
            elif v > accs[j]:
                accs[j] = v
    return [(k, accs) for k, accs in table.items()]


def _group_by(node, part, inputs, env):
    p = node.params
    src = _cat(inputs[0])
    mode = p["mode"]
    key = _kfn(p["key"])
    elem = p["elem"].fn if p.get("elem") is not None else None
    if mode == "full":
        groups = L.group_by(src, key, elem)
        if p.get("result") is not None:
            r = p["result"].fn
            return [[r(g.key, g) for g in groups]]
        return [groups]
    dec = p["decomp"]
    table: Dict[Any, list] = {}
    keys: Dict[Any, Any] = {}
    if mode in ("partial_acc", "full_acc"):
        fast = _fast_accumulate(src, key, elem, dec)
        if fast is not None:
            return [fast]
        for x in src:
            k = key(x)
            hk = L._hashable(k)
            v = elem(x) if elem else x
            accs = table.get(hk)
            if accs is None:
                table[hk] = [a.seed(v) for a in dec.aggs]
                keys[hk] = k
            else:
                for j, a in enumerate(dec.aggs):
                    accs[j] = a.accumulate(accs[j], v)
    else:  # final_acc: combine partial (key, accs) pairs
        for k, paccs in src:
            hk = L._hashable(k)
            accs = table.get(hk)
            if accs is None:
                table[hk] = list(paccs)
                keys[hk] = k
            else:
                for j, a in enumerate(dec.aggs):
                    accs[j] = a.combine(accs[j], paccs[j])
    return [[(keys[hk], accs) for hk, accs in table.items()]]


def _join(node, part, inputs, env):
    p = node.params
    outer, inner = _cat(inputs[0]), _cat(inputs[1])
    ok, ik, res = _kfn(p["outer_key"]), _kfn(p["inner_key"]), p["result"].fn
    if node.type == NodeType.GroupJoin:
        return [L.group_join(outer, inner, ok, ik, res)]
    if p.get("algorithm") == "merge":
        return [_merge_join(outer, inner, ok, ik, res, bool(p.get("descending")))]
    return [L.join(outer, inner, ok, ik, res)]


def _merge_join(outer, inner, ok, ik, res, desc):
    """Co-iterate two key-sorted inputs (DryadLinqVertex.MergeJoin :898-1030)."""
    out = []
    i = j = 0
    lt = (lambda a, b: a > b) if desc else (lambda a, b: a < b)
    while i < len(outer) and j < len(inner):
        ko, ki = ok(outer[i]), ik(inner[j])
        if lt(ko, ki):
            i += 1
        elif lt(ki, ko):
            j += 1
        else:
            j2 = j
            while j2 < len(inner) and ik(inner[j2]) == ko:
                j2 += 1
            while i < len(outer) and ok(outer[i]) == ko:
                for jj in range(j, j2):
                    out.append(res(outer[i], inner[jj]))
                i += 1
            j = j2
    return out


def _distinct(node, part, inputs, env):
    return [L.distinct(_cat(inputs[0]))]


def _set_operation(node, part, inputs, env):
    a, b = _cat(inputs[0]), _cat(inputs[1])
    op = node.params["op"]
    return [{"union": L.union, "intersect": L.intersect, "except": L.except_}[op](a, b)]


def _partition_op(node, part, inputs, env):
    p = node.params
    src = _cat(inputs[0])
    op = p["op"]
    if op == "take":
        return [L.take(src, p["n"])]
    if op == "skip":
        return [L.skip(src, p["n"])]
    f = p["fn"].fn
    if op == "take_while":
        return [L.take_while(src, f, p.get("indexed", False))]
    return [L.skip_while(src, f, p.get("indexed", False))]


def _zip(node, part, inputs, env):
    return [L.zip_(_cat(inputs[0]), _cat(inputs[1]), node.params["fn"].fn)]


def _reverse(node, part, inputs, env):
    return [L.reverse(_cat(inputs[0]))]


def _sliding_window(node, part, inputs, env):
    # inputs[1] = halo: the last w-1 records preceding this partition (neighbour exchange, C6)
    p = node.params
    w = p["window"]
    halo = _cat(inputs[1]) if len(inputs) > 1 else []
    data = halo + _cat(inputs[0])
    f = p["fn"].fn
    return [[f(list(data[i:i + w])) for i in range(0, len(data) - w + 1)]] if w > 0 else [[]]


# ---- aggregates ---------------------------------------------------------------------------------------------
_EMPTY = ("__empty__",)


def _basic_aggregate(node, part, inputs, env):
    p = node.params
    kind, stage = p["kind"], p["stage"]
    src = _cat(inputs[0])
    fn = p["fn"].fn if p.get("fn") is not None else None
    if kind == "sequence_equal":
        return [[L.sequence_equal(src, _cat(inputs[1]))]]
    if kind == "any_true":
        return [[any(src)]]
    if stage == 1:
        vals = src
        if kind == "count":
            return [[L.count(vals, fn)]]
        if kind in ("sum", "min", "max", "average"):
            vals = [fn(x) for x in vals] if fn else vals
            if kind == "sum":
                return [[L.sum_(vals)]]
            if not vals:
                return [[]]
            if kind == "min":
                return [[min(vals)]]
            if kind == "max":
                return [[max(vals)]]
            s = 0
            for v in vals:
                s = L.checked_add(s, v) if isinstance(v, int) else s + v
            return [[(s, len(vals))]]
        if kind == "any":
            return [[L.any_(vals, fn)]]
        if kind == "all":
            return [[L.all_(vals, fn)]]
        m = [x for x in vals if fn is None or fn(x)]
        if kind == "first":
            return [m[:1]]
        if kind == "last":
            return [m[-1:]]
        if kind == "single":
            return [m[:2]]
        raise DryadLinqException(ErrorCode.AggregateOperatorNotSupported, kind)
    # stage 2: combine partials (one value per non-empty partition, in port order)
    if kind == "count" or kind == "sum":
        return [[L.sum_(src)]]
    if kind == "any":
        return [[any(src)]]
    if kind == "all":
        return [[all(src)]]
    try:
        if kind == "min":
            return [[L.min_(src)]]
        if kind == "max":
            return [[L.max_(src)]]
        if kind == "average":
            if not src:
                raise ValueError("Sequence contains no elements (Average)")
            s, c = 0, 0
            for ps, pc in src:
                s = L.checked_add(s, ps) if isinstance(ps, int) else s + ps
                c += pc
            return [[s / c]]
        if kind == "first":
            return [[L.first(src, None, p["default"])]]
        if kind == "last":
            return [[L.last(src, None, p["default"])]]
        if kind == "single":
            return [[L.single(src, None, p["default"])]]
    except ValueError as ex:
        code = {"min": ErrorCode.MinNoElements, "max": ErrorCode.MaxNoElements,
                "average": ErrorCode.AverageNoElements, "first": ErrorCode.FirstNoElementsFirst,
                "last": ErrorCode.LastNoElements, "single": ErrorCode.SingleNoElements}[kind]
        if kind == "single" and "more than one" in str(ex):
            code = ErrorCode.SingleMoreThanOneElement
        raise DryadLinqException(code, str(ex)) from ex
    raise DryadLinqException(ErrorCode.AggregateOperatorNotSupported, kind)


# ---- internal vertices of a run-time aggregation tree -----------------------------------------------------------
_COMMUTATIVE_KINDS = ("count", "sum", "min", "max", "any", "all", "average", "any_true")


def combiner_of(node):
    """The member of an N:1 consumer that combines per-partition partials — stage 2 of an Aggregate or
    BasicAggregate directly behind the consumer's Merge — or None.  With it: (member, commutative)."""
    members = node.members if node.type == NodeType.Super else [node]
    for m in members:
        if m.type == NodeType.Merge and m.params.get("sort_key") is None:
            continue
        if m.type == NodeType.BasicAggregate and m.params.get("stage") == 2 and m.params["kind"] != "sequence_equal":
            return m, m.params["kind"] in _COMMUTATIVE_KINDS
        if m.type == NodeType.Aggregate and m.params.get("stage") == 2:
            return m, False                     # associative, not known to be commutative: contiguous groups only
        if m.type == NodeType.GroupBy and m.params.get("mode") == "final_acc":
            return m, True                      # (key, accumulators) pairs in, the same shape out
        return None
    return None


def internal_combine(node, payloads, env) -> List:
    """What an internal vertex of the aggregation tree computes (the reference copies the consumer
    stage's vertex as the internal vertex, GraphManager/stagemanager/DrDynamicAggregateManager.h:
    24-179): it folds SOME of the consumer's partial inputs into ONE partial of the same shape, so the
    consumer — or another internal vertex — can take it in their place."""
    found = combiner_of(node)
    if found is None:
        raise DryadLinqException(ErrorCode.Internal, f"{node.name} has no combiner for an aggregation tree")
    m = found[0]
    src = _cat(payloads)
    p = m.params
    if m.type == NodeType.GroupBy:
        return _group_by(m, 0, [[src]], env)[0]
    if m.type == NodeType.Aggregate:
        return [L.aggregate(src, p["fn"].fn)] if src else []
    kind = p["kind"]
    if kind in ("count", "sum"):
        return [L.sum_(src)]
    if kind in ("any", "any_true"):
        return [any(src)]
    if kind == "all":
        return [all(src)]
    if not src:
        return []
    if kind == "min":
        return [min(src)]
    if kind == "max":
        return [max(src)]
    if kind == "average":
        s, c = 0, 0
        for ps, pc in src:
            s = L.checked_add(s, ps) if isinstance(ps, int) else s + ps
            c += pc
        return [(s, c)]
    if kind == "first":
        return src[:1]
    if kind == "last":
        return src[-1:]
    if kind == "single":
        return src[:2]
    raise DryadLinqException(ErrorCode.AggregateOperatorNotSupported, kind)


def _contains(node, part, inputs, env):
    return [[L.contains(_cat(inputs[0]), node.params["value"])]]


def _aggregate(node, part, inputs, env):
    p = node.params
    src = _cat(inputs[0])
    f = p["fn"].fn
    stage = p["stage"]
    if stage in (1, 2):
        if not src:
            if stage == 2:
                raise DryadLinqException(ErrorCode.AggregateNoElements, "Sequence contains no elements (Aggregate)")
            return [[]]
        return [[L.aggregate(src, f)]]
    seed = p["seed_fn"]() if p.get("seed_fn") else p.get("seed")
    try:
        return [[L.aggregate(src, f, seed, p["has_seed"], p["result"].fn if p.get("result") else None)]]
    except ValueError as ex:
        raise DryadLinqException(ErrorCode.AggregateNoElements, str(ex)) from ex


# ---- apply / fork -------------------------------------------------------------------------------------------------
def _apply(node, part, inputs, env):
    p = node.params
    f = p["fn"].fn
    a = _cat(inputs[0])
    try:
        if p.get("with_index"):
            return [list(f(a, part))]
        if len(inputs) > 1:
            return [list(f(a, _cat(inputs[1])))]
        return [list(f(a))]
    except DryadLinqException:
        raise
    except Exception as ex:
        raise DryadLinqException(ErrorCode.FailureInUserApplyFunction, f"{type(ex).__name__}: {ex}", ex) from ex


def _fork(node, part, inputs, env):
    p = node.params
    src = _cat(inputs[0])
    if p["per_partition"]:
        outs = p["fn"].fn(src)
        return [list(o) for o in outs]
    return [list(o) for o in L.fork(src, p["fn"].fn, p["n"])]


_DISPATCH = {
    NodeType.InputTable: _input, NodeType.OutputTable: _output, NodeType.Dummy: _identity,
    NodeType.Merge: _merge, NodeType.Tee: _tee, NodeType.Concat: _concat, NodeType.ForkBranch: _fork_branch,
    NodeType.Select: _select, NodeType.Where: _where, NodeType.SelectMany: _select_many,
    NodeType.HashPartition: _hash_partition, NodeType.RangePartition: _range_partition,
    NodeType.Sampler: _sampler, NodeType.RangeSampler: _range_sampler, NodeType.Offsets: _offsets,
    NodeType.OrderBy: _order_by, NodeType.GroupBy: _group_by, NodeType.Join: _join, NodeType.GroupJoin: _join,
    NodeType.Distinct: _distinct, NodeType.SetOperation: _set_operation, NodeType.PartitionOp: _partition_op,
    NodeType.Zip: _zip, NodeType.Reverse: _reverse, NodeType.SlidingWindow: _sliding_window,
    NodeType.BasicAggregate: _basic_aggregate, NodeType.Contains: _contains, NodeType.Aggregate: _aggregate,
    NodeType.Apply: _apply, NodeType.Fork: _fork,
}
