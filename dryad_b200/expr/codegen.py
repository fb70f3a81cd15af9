"""Per-pipeline CUDA code generation with NVRTC — the counterpart of DryadLinqCodeGen, which emits
and compiles one vertex method per fused pipeline (LinqToDryad/DryadLinqCodeGen.cs:1910-2311).

A fused stage's projections and predicate are rendered as straight-line CUDA C (one grid-stride
kernel, each input column loaded once into a register, each output stored once), compiled for
sm_100a with NVRTC and cached by source text.  The generic register-program interpreter
(`csrc/kernels/stream.cu`, `compile.run_on_gpu`) stays as the fallback when NVRTC is unavailable.
"""
from __future__ import annotations

import ctypes
import struct
from typing import Dict, Mapping, Optional, Tuple

import torch

from .ir import Expr, TraceError, retype, wrap

_CT = {torch.int64: "long long", torch.float64: "double", torch.int32: "int", torch.float32: "float",
       torch.uint8: "unsigned char", torch.int16: "short", torch.bool: "unsigned char"}
_OUT = {"i": ("long long", torch.int64), "f": ("double", torch.float64), "b": ("unsigned char", torch.uint8)}
_BIN = {"add": "+", "sub": "-", "mul": "*", "eq": "==", "ne": "!=", "lt": "<", "le": "<=", "gt": ">", "ge": ">=",
        "band": "&", "bor": "|", "bxor": "^", "shl": "<<", "shr": ">>"}

_PRELUDE = r"""
typedef long long i64; typedef unsigned long long u64;
__device__ __forceinline__ u64 dryad_mix64(u64 x) {
  x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ULL; x ^= x >> 27; x *= 0x94d049bb133111ebULL; x ^= x >> 31; return x; }
// Python's // and %: floored quotient, remainder with the divisor's sign (0 for a zero divisor)
__device__ __forceinline__ i64 dryad_div(i64 a, i64 b) {
  if (!b) return 0;
  i64 q = a / b, r = a % b;
  return (r != 0 && ((r < 0) != (b < 0))) ? q - 1 : q; }
__device__ __forceinline__ i64 dryad_mod(i64 a, i64 b) {
  if (!b) return 0;
  i64 r = a % b;
  return (r != 0 && ((r < 0) != (b < 0))) ? r + b : r; }
__device__ __forceinline__ double dryad_fmod(double a, double b) {
  double r = fmod(a, b);
  return (r != 0.0 && ((r < 0.0) != (b < 0.0))) ? r + b : r; }
// read-only probe of a finished join table (4-key buckets, csrc/kernels/hash_table.cu find_slot)
template <bool WANT_SLOT>
__device__ __forceinline__ i64 dryad_probe(const u64* __restrict__ keys, const int* __restrict__ head, u64 cap, u64 key) {
  const u64 nb = cap >> 2;
  u64 b = __umul64hi(dryad_mix64(key), nb);
  for (u64 p = 0; p < nb; ++p) {
    u64 k0, k1, k2, k3;
    // two 128-bit loads of one 32-byte sector (the NVRTC in this image predates 256-bit ld)
    asm volatile("ld.global.nc.v2.u64 {%0, %1}, [%2];" : "=l"(k0), "=l"(k1) : "l"(keys + b * 4));
    asm volatile("ld.global.nc.v2.u64 {%0, %1}, [%2];" : "=l"(k2), "=l"(k3) : "l"(keys + b * 4 + 2));
    const u64 k[4] = {k0, k1, k2, k3};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (k[j] == key) return WANT_SLOT ? (i64)(b * 4 + j) : (i64)__ldg(head + b * 4 + j);
      if (k[j] == ~0ULL) return -1;
    }
    if (++b == nb) b = 0;
  }
  return -1;
}
template <typename T> __device__ __forceinline__ T dryad_warp_red(T v, int op) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    const T o = __shfl_xor_sync(0xffffffffu, v, d);
    v = op == 0 ? v + o : op == 1 ? (o < v ? o : v) : (o > v ? o : v);
  }
  return v;
}
"""

_cache: Dict[str, Tuple[object, object]] = {}
_available: Optional[bool] = None


class CodegenUnavailable(Exception):
    pass


def _ctype_of_kind(k: str) -> str:
    return {"i": "i64", "f": "double", "b": "int"}[k]


class _Ctx:
    """Per-kernel emission state: hoisted statements (probes, gather indices) with CSE."""

    def __init__(self, tables: Mapping[str, int], arrays: Mapping[str, Tuple[int, str]]):
        self.tables, self.arrays = tables, arrays
        self.stmts: list = []
        self.memo: Dict[tuple, str] = {}

    def hoist(self, key: tuple, ctype: str, code: str) -> str:
        v = self.memo.get(key)
        if v is None:
            v = f"h{len(self.memo)}"
            self.memo[key] = v
            self.stmts.append(f"    const {ctype} {v} = {code};")
        return v


def emit(e: Expr, cols: Mapping[str, str], ctx: Optional[_Ctx] = None) -> str:
    """C expression for e; `cols` maps column name -> local variable name."""
    op, k = e.op, e.kind
    if k == "o":
        raise TraceError("opaque expression has no GPU code")
    if op == "col":
        return cols[e.args[0]]
    if op == "lit":
        v = e.args[0]
        if k == "f":
            return f"__longlong_as_double({struct.unpack('<q', struct.pack('<d', float(v)))[0]}LL)"
        return f"((i64){int(v)}LL)" if k == "i" else ("1" if v else "0")
    if op == "rowid":
        return "i"
    if op in ("probe", "probe_slot"):
        if ctx is None or e.args[0] not in ctx.tables:
            raise TraceError(f"probe of unknown table {e.args[0]!r}")
        t = ctx.tables[e.args[0]]
        key = e.args[1]
        kc = emit(key, cols, ctx)
        kc = f"(u64)__double_as_longlong({kc})" if key.kind == "f" else f"(u64)({kc})"
        want = "true" if op == "probe_slot" else "false"
        return ctx.hoist((op, t, kc), "i64", f"dryad_probe<{want}>(tk{t}, th{t}, tc{t}, {kc})")
    if op == "gather":
        if ctx is None or e.args[0] not in ctx.arrays:
            raise TraceError(f"gather from unknown array {e.args[0]!r}")
        a, akind = ctx.arrays[e.args[0]]
        idx = ctx.hoist(("idx", emit(e.args[1], cols, ctx)), "i64", emit(e.args[1], cols, ctx))
        ct = "double" if akind == "f" else "i64"
        return ctx.hoist(("gather", a, idx), ct, f"({idx} >= 0 ? ({ct})__ldg(arr{a} + {idx}) : ({ct})0)")
    A = [emit(a, cols, ctx) for a in e.args if isinstance(a, Expr)]
    if op in _BIN:
        if op in ("eq", "ne", "lt", "le", "gt", "ge"):
            return f"(({A[0]}) {_BIN[op]} ({A[1]}) ? 1 : 0)"
        if op in ("shl", "shr"):
            return f"(({A[0]}) {_BIN[op]} (({A[1]}) & 63))"
        return f"(({A[0]}) {_BIN[op]} ({A[1]}))"
    if op == "div":
        return f"(({A[0]}) / ({A[1]}))" if k == "f" else f"dryad_div({A[0]}, {A[1]})"
    if op == "mod":
        if k == "f":
            return f"dryad_fmod({A[0]}, {A[1]})"
        return f"dryad_mod({A[0]}, {A[1]})"
    if op == "and":
        return f"((({A[0]}) != 0) & (({A[1]}) != 0))"
    if op == "or":
        return f"((({A[0]}) != 0) | (({A[1]}) != 0))"
    if op == "not":
        return f"(({A[0]}) == 0 ? 1 : 0)"
    if op == "bnot":
        return f"(~({A[0]}))"
    if op == "neg":
        return f"(-({A[0]}))"
    if op == "abs":
        return f"fabs({A[0]})" if k == "f" else f"(({A[0]}) < 0 ? -({A[0]}) : ({A[0]}))"
    if op == "min":
        return f"fmin({A[0]}, {A[1]})" if k == "f" else f"(({A[0]}) < ({A[1]}) ? ({A[0]}) : ({A[1]}))"
    if op == "max":
        return f"fmax({A[0]}, {A[1]})" if k == "f" else f"(({A[0]}) > ({A[1]}) ? ({A[0]}) : ({A[1]}))"
    if op == "cast":
        if k == "f":
            return f"((double)({A[0]}))"
        if k == "i":
            return f"((i64)({A[0]}))"
        return f"(({A[0]}) != 0 ? 1 : 0)"
    if op == "select":
        return f"(({A[0]}) ? ({A[1]}) : ({A[2]}))"
    if op == "hash":
        src = e.args[0]
        x = f"(u64)__double_as_longlong({A[0]})" if src.kind == "f" else f"(u64)({A[0]})"
        return f"((i64)dryad_mix64({x}))"
    if op == "sqrt":
        return f"sqrt({A[0]})"
    if op == "floor":
        return f"floor({A[0]})"
    raise TraceError(f"op {op!r} has no GPU code")


_RED = {"sum": 0, "min": 1, "max": 2}
_RED_ID = {("sum", "i"): "0", ("sum", "f"): "0.0", ("min", "i"): "0x7fffffffffffffffLL", ("max", "i"): "(-0x7fffffffffffffffLL - 1)",
           ("min", "f"): "__longlong_as_double(0x7ff0000000000000LL)", ("max", "f"): "__longlong_as_double(0xfff0000000000000LL)"}


def _source(outputs: Mapping[str, Expr], predicate: Optional[Expr], in_cols, in_dtypes,
            tables=(), arrays=(), reduce: Optional[Mapping[str, str]] = None, count_flags: bool = False) -> str:
    """tables: names; arrays: (name, dtype) pairs; reduce: output name -> 'sum'|'min'|'max' (those
    outputs become per-block partials instead of columns); count_flags: also reduce the predicate
    to a per-block count instead of storing a flag per row."""
    reduce = dict(reduce or {})
    params, loads, names = [], [], {}
    for j, (c, dt) in enumerate(zip(in_cols, in_dtypes)):
        params.append(f"const {_CT[dt]}* __restrict__ in{j}")
        kind = "double" if dt in (torch.float32, torch.float64) else "i64"
        loads.append(f"    const {kind} c{j} = ({kind})in{j}[i];")
        names[c] = f"c{j}"
    ctx = _Ctx({t: j for j, t in enumerate(tables)},
               {a: (j, "f" if dt in (torch.float32, torch.float64) else "i") for j, (a, dt) in enumerate(arrays)})
    for j, _ in enumerate(tables):
        params += [f"const u64* __restrict__ tk{j}", f"const int* __restrict__ th{j}", f"u64 tc{j}"]
    for j, (_, dt) in enumerate(arrays):
        params.append(f"const {_CT[dt]}* __restrict__ arr{j}")
    body, accs, fin = [], [], []
    for j, (name, e) in enumerate(outputs.items()):
        code = emit(e, names, ctx)
        if name in reduce:
            if e.kind not in ("i", "f"):
                raise TraceError("only numeric outputs can be reduced")
            ct = "double" if e.kind == "f" else "i64"
            opn = reduce[name]
            params.append(f"{ct}* __restrict__ out{j}")
            accs.append(f"  {ct} acc{j} = {_RED_ID[(opn, e.kind)]};")
            upd = {"sum": f"acc{j} + v{j}", "min": f"(v{j} < acc{j} ? v{j} : acc{j})", "max": f"(v{j} > acc{j} ? v{j} : acc{j})"}[opn]
            body.append(f"    {{ const {ct} v{j} = ({ct})({code}); acc{j} = {upd}; }}")
            fin.append((f"acc{j}", f"out{j}", ct, _RED[opn], _RED_ID[(opn, e.kind)]))
        else:
            params.append(f"{_OUT[e.kind][0]}* __restrict__ out{j}")
            body.append(f"    out{j}[i] = ({_OUT[e.kind][0]})({code});")
    params.append("unsigned char* __restrict__ flags" if not count_flags else "i64* __restrict__ flags")
    params.append("i64 n")
    if predicate is not None:
        pc = emit(predicate, names, ctx)
        if count_flags:
            accs.append("  i64 accp = 0;")
            body.append(f"    accp += (({pc}) != 0);")
            fin.append(("accp", "flags", "i64", 0, "0"))
        else:
            body.append(f"    flags[i] = (unsigned char)(({pc}) != 0);")
    tail = []
    if fin:
        tail.append("  __shared__ double red_smem[8];")
        for acc, dst, ct, opn, ident in fin:
            sm = f"reinterpret_cast<{ct}*>(red_smem)"
            tail += [f"  {acc} = dryad_warp_red<{ct}>({acc}, {opn});",
                     f"  if ((threadIdx.x & 31) == 0) {sm}[threadIdx.x >> 5] = {acc};",
                     "  __syncthreads();",
                     f"  if (threadIdx.x < 32) {{ {ct} t = threadIdx.x < (blockDim.x >> 5) ? {sm}[threadIdx.x] : ({ct}){ident};"
                     f" t = dryad_warp_red<{ct}>(t, {opn}); if (threadIdx.x == 0) {dst}[blockIdx.x] = t; }}",
                     "  __syncthreads();"]
    return (_PRELUDE + 'extern "C" __global__ void __launch_bounds__(256) dryad_fused(' + ", ".join(params) + ") {\n"
            + "\n".join(accs) + "\n"
            "  i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x;\n  const i64 stride = gridDim.x * (i64)blockDim.x;\n"
            "  for (; i < n; i += stride) {\n" + "\n".join(loads + ctx.stmts + body) + "\n  }\n" + "\n".join(tail) + "\n}\n")


def _compile(src: str):
    from cuda.bindings import driver, nvrtc
    hit = _cache.get(src)
    if hit is not None:
        return hit
    err, prog = nvrtc.nvrtcCreateProgram(src.encode(), b"dryad_fused.cu", 0, [], [])
    opts = [b"--gpu-architecture=sm_100a", b"--std=c++17", b"-lineinfo"]
    (err,) = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    if int(err) != 0:
        _, n = nvrtc.nvrtcGetProgramLogSize(prog)
        log = b" " * n
        nvrtc.nvrtcGetProgramLog(prog, log)
        raise CodegenUnavailable("nvrtc: " + log.decode(errors="replace")[-800:])
    _, n = nvrtc.nvrtcGetCUBINSize(prog)
    cubin = b" " * n
    nvrtc.nvrtcGetCUBIN(prog, cubin)
    err, mod = driver.cuModuleLoadData(cubin)
    if int(err) != 0:
        raise CodegenUnavailable(f"cuModuleLoadData failed: {err}")
    err, fn = driver.cuModuleGetFunction(mod, b"dryad_fused")
    if int(err) != 0:
        raise CodegenUnavailable(f"cuModuleGetFunction failed: {err}")
    _cache[src] = (mod, fn)
    return mod, fn


def available() -> bool:
    global _available
    if _available is None:
        try:
            from cuda.bindings import driver, nvrtc  # noqa: F401
            _available = True
        except Exception:  # noqa: BLE001
            _available = False
    return _available


def run_fused(outputs: Mapping[str, Expr], cols: Mapping[str, torch.Tensor], n: int,
              predicate: Optional[Expr] = None, *, tables: Optional[Mapping[str, object]] = None,
              arrays: Optional[Mapping[str, torch.Tensor]] = None,
              reduce: Optional[Mapping[str, str]] = None, count_flags: bool = False):
    """Evaluate projections (+ predicate) with a generated, NVRTC-compiled kernel.
    Returns (out_cols, flags) like compile.run_on_gpu; raises CodegenUnavailable to request the
    interpreter fallback.

    tables: name -> object with `.keys`, `.head`, `.cap` (a finished unique-key join table) for
    `probe(name, key)` nodes; arrays: name -> 1-D tensor for `gather(name, index)` nodes.
    reduce: output name -> 'sum' | 'min' | 'max': that output is returned as a 0-d tensor (per-block
    partials folded by the deterministic K.reduce) instead of a column; count_flags: return the
    number of rows passing the predicate (0-d int64) instead of a flag per row."""
    if not available():
        raise CodegenUnavailable("cuda-python / nvrtc not importable")
    from cuda.bindings import driver
    from ..ops import core as K
    from .compile import kind_of_dtype
    tables = dict(tables or {})
    arrays = dict(arrays or {})
    reduce = dict(reduce or {})
    dts = {k: v.dtype for k, v in cols.items() if isinstance(v, torch.Tensor) and v.dim() == 1}
    kinds = {k: kind_of_dtype(dt) for k, dt in dts.items()}
    outs_e = {k: retype(wrap(e), kinds) for k, e in outputs.items()}
    pred = retype(predicate, kinds) if predicate is not None else None
    needed = set()
    for e in outs_e.values():
        needed |= e.columns()
    if pred is not None:
        needed |= pred.columns()
    in_cols = sorted(needed)
    for c in in_cols:
        if c not in dts or dts[c] not in _CT:
            raise CodegenUnavailable(f"column {c!r} unsupported")
    tensors = [cols[c] if cols[c].is_contiguous() else cols[c].contiguous() for c in in_cols]
    tnames = sorted(tables)
    anames = sorted(arrays)
    for a in anames:
        if arrays[a].dtype not in _CT or not arrays[a].is_contiguous():
            raise CodegenUnavailable(f"array {a!r} unsupported")
    try:
        src = _source(outs_e, pred, in_cols, [t.dtype for t in tensors], tnames,
                      [(a, arrays[a].dtype) for a in anames], reduce, count_flags)
    except TraceError as ex:
        raise CodegenUnavailable(str(ex)) from ex
    _, fn = _compile(src)
    dev = tensors[0].device if tensors else torch.device("cuda")
    grid = max(1, min((n + 256 * 4 - 1) // (256 * 4), 148 * 16))
    result: Dict[str, torch.Tensor] = {}
    out_t = []
    for name, e in outs_e.items():
        if name in reduce:
            t = torch.empty(grid, dtype=torch.float64 if e.kind == "f" else torch.int64, device=dev)
        else:
            t = torch.empty(n, dtype=_OUT[e.kind][1], device=dev)
        result[name] = t
        out_t.append(t)
    flags = None
    if pred is not None:
        flags = torch.empty(grid, dtype=torch.int64, device=dev) if count_flags else torch.empty(n, dtype=torch.uint8, device=dev)
    if n > 0:
        vals = [t.data_ptr() for t in tensors]
        types = [ctypes.c_void_p] * len(tensors)
        for tn in tnames:
            tb = tables[tn]
            vals += [tb.keys.data_ptr(), tb.head.data_ptr(), int(tb.cap)]
            types += [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_ulonglong]
        vals += [arrays[a].data_ptr() for a in anames] + [t.data_ptr() for t in out_t]
        types += [ctypes.c_void_p] * (len(anames) + len(out_t))
        vals += [flags.data_ptr() if flags is not None else 0, n]
        types += [ctypes.c_void_p, ctypes.c_longlong]
        stream = torch.cuda.current_stream().cuda_stream
        (err,) = driver.cuLaunchKernel(fn, grid, 1, 1, 256, 1, 1, 0, stream, (tuple(vals), tuple(types)), 0)
        if int(err) != 0:
            raise CodegenUnavailable(f"cuLaunchKernel failed: {err}")
        K._count()
    for name, e in outs_e.items():
        if name in reduce:
            ident = {"sum": 0, "min": None, "max": None}[reduce[name]]
            result[name] = K.reduce(result[name], reduce[name]) if n > 0 else torch.tensor(
                ident if ident is not None else (float("inf") if reduce[name] == "min" else float("-inf")), device=dev)
        elif e.kind == "b":
            result[name] = result[name].view(torch.bool)
    if count_flags and flags is not None:
        flags = K.reduce(flags, "sum") if n > 0 else torch.zeros((), dtype=torch.int64, device=dev)
    return result, flags
