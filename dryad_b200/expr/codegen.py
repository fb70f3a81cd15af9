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
__device__ __forceinline__ i64 dryad_div(i64 a, i64 b) { return b ? a / b : 0; }
__device__ __forceinline__ i64 dryad_mod(i64 a, i64 b) { return b ? a % b : 0; }
"""

_cache: Dict[str, Tuple[object, object]] = {}
_available: Optional[bool] = None


class CodegenUnavailable(Exception):
    pass


def _ctype_of_kind(k: str) -> str:
    return {"i": "i64", "f": "double", "b": "int"}[k]


def emit(e: Expr, cols: Mapping[str, str]) -> str:
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
    A = [emit(a, cols) for a in e.args if isinstance(a, Expr)]
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
            return f"fmod({A[0]}, {A[1]})"
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
        src = e.args[0]
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


def _source(outputs: Mapping[str, Expr], predicate: Optional[Expr], in_cols, in_dtypes) -> str:
    params, loads, names = [], [], {}
    for j, (c, dt) in enumerate(zip(in_cols, in_dtypes)):
        params.append(f"const {_CT[dt]}* __restrict__ in{j}")
        kind = "double" if dt in (torch.float32, torch.float64) else "i64"
        loads.append(f"    const {kind} c{j} = ({kind})in{j}[i];")
        names[c] = f"c{j}"
    body = []
    for j, (name, e) in enumerate(outputs.items()):
        params.append(f"{_OUT[e.kind][0]}* __restrict__ out{j}")
        body.append(f"    out{j}[i] = ({_OUT[e.kind][0]})({emit(e, names)});")
    params.append("unsigned char* __restrict__ flags")
    params.append("i64 n")
    if predicate is not None:
        body.append(f"    flags[i] = (unsigned char)(({emit(predicate, names)}) != 0);")
    return (_PRELUDE + 'extern "C" __global__ void __launch_bounds__(256) dryad_fused(' + ", ".join(params) + ") {\n"
            "  i64 i = blockIdx.x * (i64)blockDim.x + threadIdx.x;\n  const i64 stride = gridDim.x * (i64)blockDim.x;\n"
            "  for (; i < n; i += stride) {\n" + "\n".join(loads + body) + "\n  }\n}\n")


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
              predicate: Optional[Expr] = None):
    """Evaluate projections (+ predicate) with a generated, NVRTC-compiled kernel.
    Returns (out_cols, flags) like compile.run_on_gpu; raises CodegenUnavailable to request the
    interpreter fallback."""
    if not available():
        raise CodegenUnavailable("cuda-python / nvrtc not importable")
    from cuda.bindings import driver
    from ..ops import core as K
    from .compile import kind_of_dtype
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
    try:
        src = _source(outs_e, pred, in_cols, [t.dtype for t in tensors])
    except TraceError as ex:
        raise CodegenUnavailable(str(ex)) from ex
    _, fn = _compile(src)
    dev = tensors[0].device if tensors else torch.device("cuda")
    result: Dict[str, torch.Tensor] = {}
    out_t = []
    for name, e in outs_e.items():
        t = torch.empty(n, dtype=_OUT[e.kind][1], device=dev)
        result[name] = t
        out_t.append(t)
    flags = torch.empty(n, dtype=torch.uint8, device=dev) if pred is not None else None
    if n > 0:
        vals = [t.data_ptr() for t in tensors] + [t.data_ptr() for t in out_t] + [flags.data_ptr() if flags is not None else 0, n]
        types = [ctypes.c_void_p] * (len(tensors) + len(out_t) + 1) + [ctypes.c_longlong]
        grid = max(1, min((n + 256 * 4 - 1) // (256 * 4), 148 * 16))
        stream = torch.cuda.current_stream().cuda_stream
        (err,) = driver.cuLaunchKernel(fn, grid, 1, 1, 256, 1, 1, 0, stream, (tuple(vals), tuple(types)), 0)
        if int(err) != 0:
            raise CodegenUnavailable(f"cuLaunchKernel failed: {err}")
        K._count()
    for name, e in outs_e.items():
        if e.kind == "b":
            result[name] = result[name].view(torch.bool)
    return result, flags
