"""Lower Expr trees to the GPU register program run by `dryad_expr_run` (csrc/kernels/stream.cu).

This is the role DryadLinqCodeGen plays for the reference (one generated method per fused
pipeline, LinqToDryad/DryadLinqCodeGen.cs:1910-2138): a fused stage's projections and predicate
become ONE kernel launch that reads each input column once.
"""
from __future__ import annotations

import struct
from typing import Dict, List, Mapping, Optional, Tuple

import torch

from .. import _native as N
from .ir import Expr, TraceError, cast, wrap

# opcode numbering must match the enum in include/dryad_b200.h
_OPS = ["LOAD", "CONST", "ROWID", "ADD_I", "SUB_I", "MUL_I", "DIV_I", "MOD_I", "NEG_I", "ADD_F",
        "SUB_F", "MUL_F", "DIV_F", "NEG_F", "EQ_I", "NE_I", "LT_I", "LE_I", "GT_I", "GE_I", "EQ_F",
        "NE_F", "LT_F", "LE_F", "GT_F", "GE_F", "AND", "OR", "NOT", "BAND", "BOR", "BXOR", "BNOT",
        "SHL", "SHR", "I2F", "F2I", "SELECT", "HASH", "ABS_I", "ABS_F", "MIN_I", "MAX_I", "MIN_F",
        "MAX_F", "SQRT_F", "FLOOR_F", "MOV", "STORE"]
OP = {n: i for i, n in enumerate(_OPS)}

_DT = {torch.int64: 0, torch.float64: 1, torch.int32: 2, torch.float32: 3, torch.uint8: 4,
       torch.int16: 5, torch.bool: 4}
_OUT_DT = {"i": (torch.int64, 0), "f": (torch.float64, 1), "b": (torch.uint8, 4)}


class CompileError(Exception):
    pass


def kind_of_dtype(dt: torch.dtype) -> str:
    if dt in (torch.float32, torch.float64, torch.float16, torch.bfloat16):
        return "f"
    if dt == torch.bool:
        return "b"
    return "i"


class _Builder:
    def __init__(self, col_index: Mapping[str, int]):
        self.col_index = dict(col_index)
        self.ins: List[Tuple[int, int, int, int, int]] = []
        self.free = list(range(15, -1, -1))
        self.memo: Dict[tuple, int] = {}
        self.pinned: set = set()

    def alloc(self) -> int:
        if not self.free:
            raise CompileError("expression needs more than 16 registers")
        return self.free.pop()

    def release(self, r: int) -> None:
        if r not in self.pinned and r not in self.free:
            self.free.append(r)

    def emit(self, op: str, dst: int, a: int = 0, b: int = 0, imm: int = 0) -> None:
        if len(self.ins) >= N.EXPR_MAX_INSTR:
            raise CompileError("expression program too long")
        self.ins.append((OP[op], dst, a, b, imm))

    def gen(self, e: Expr) -> int:
        """Emit code for e; returns the register holding it (caller releases)."""
        op, k = e.op, e.kind
        if op == "col":
            name = e.args[0]
            if name not in self.col_index:
                raise CompileError(f"unknown column {name!r}")
            r = self.alloc()
            self.emit("LOAD", r, self.col_index[name])
            return r
        if op == "lit":
            v = e.args[0]
            if k == "o":
                raise CompileError("opaque literal")
            imm = struct.unpack("<q", struct.pack("<d", float(v)))[0] if k == "f" else int(v)
            r = self.alloc()
            self.emit("CONST", r, imm=imm)
            return r
        if op == "rowid":
            r = self.alloc()
            self.emit("ROWID", r)
            return r
        if op == "select":
            c, a, b = (self.gen(x) for x in e.args)
            r = self.alloc()
            self.emit("SELECT", r, c, a, imm=b)
            for x in (c, a, b):
                self.release(x)
            return r
        if op == "cast":
            src = e.args[0]
            a = self.gen(src)
            if k == src.kind or {k, src.kind} == {"i", "b"} and k == "i":
                return a
            r = self.alloc()
            if k == "f":
                self.emit("I2F", r, a)
            elif k == "i":
                self.emit("F2I", r, a)
            elif k == "b":
                z = self.alloc()
                self.emit("CONST", z, imm=0)
                self.emit("NE_F" if src.kind == "f" else "NE_I", r, a, z)
                self.release(z)
            else:
                raise CompileError("cast to opaque")
            self.release(a)
            return r
        args = [x for x in e.args if isinstance(x, Expr)]
        ak = args[0].kind if args else k
        fl = "F" if (ak == "f" or (len(args) > 1 and args[1].kind == "f")) else "I"
        table = {
            "add": f"ADD_{fl}", "sub": f"SUB_{fl}", "mul": f"MUL_{fl}", "div": f"DIV_{fl}",
            "mod": "MOD_I", "eq": f"EQ_{fl}", "ne": f"NE_{fl}", "lt": f"LT_{fl}", "le": f"LE_{fl}",
            "gt": f"GT_{fl}", "ge": f"GE_{fl}", "and": "AND", "or": "OR", "band": "BAND",
            "bor": "BOR", "bxor": "BXOR", "shl": "SHL", "shr": "SHR", "min": f"MIN_{fl}",
            "max": f"MAX_{fl}",
        }
        unary = {"neg": f"NEG_{fl}", "abs": f"ABS_{fl}", "not": "NOT", "bnot": "BNOT", "hash": "HASH",
                 "sqrt": "SQRT_F", "floor": "FLOOR_F"}
        if op in table:
            if op == "mod" and fl == "F":
                raise CompileError("float remainder is host-only")
            x, y = args
            if fl == "F":
                x, y = cast(x, "f"), cast(y, "f")
            a = self.gen(x)
            b = self.gen(y)
            r = self.alloc()
            self.emit(table[op], r, a, b)
            self.release(a)
            self.release(b)
            return r
        if op in unary:
            a = self.gen(args[0])
            r = self.alloc()
            self.emit(unary[op], r, a)
            self.release(a)
            return r
        raise CompileError(f"op {op!r} has no GPU lowering")


def compile_program(outputs: Mapping[str, Expr], predicate: Optional[Expr], col_dtypes: Mapping[str, torch.dtype]):
    """Returns (instruction tuples, input column order, output (name, kind) order, result_reg)."""
    needed = set()
    for e in outputs.values():
        needed |= wrap(e).columns()
    if predicate is not None:
        needed |= predicate.columns()
    in_cols = sorted(needed)
    if len(in_cols) > N.EXPR_MAX_COLS or len(outputs) > N.EXPR_MAX_COLS:
        raise CompileError("too many columns for one fused program")
    for c in in_cols:
        if c not in col_dtypes:
            raise CompileError(f"unknown column {c!r}")
        if col_dtypes[c] not in _DT:
            raise CompileError(f"column {c!r} has unsupported dtype {col_dtypes[c]}")
    b = _Builder({c: i for i, c in enumerate(in_cols)})
    out_order = []
    for j, (name, e) in enumerate(outputs.items()):
        e = wrap(e)
        if e.kind == "o":
            raise CompileError("opaque output")
        r = b.gen(e)
        b.emit("STORE", 0, r, j)
        b.release(r)
        out_order.append((name, e.kind))
    result_reg = 0
    if predicate is not None:
        r = b.gen(predicate)
        result_reg = r
    return b.ins, in_cols, out_order, result_reg


def run_on_gpu(outputs: Mapping[str, Expr], cols: Mapping[str, torch.Tensor], n: int,
               predicate: Optional[Expr] = None, *, tables=None, arrays=None, reduce=None,
               count_flags: bool = False):
    """Evaluate projections (+ optional predicate) in ONE kernel.  Returns (out_cols, flags).
    Preferred path: a kernel generated for this exact pipeline and compiled with NVRTC
    (expr/codegen.py); fallback: the generic register-program interpreter kernel.
    `tables` / `arrays` / `reduce` / `count_flags` (fused join probes, side-array gathers and
    in-kernel reductions, see codegen.run_fused) exist only in generated code."""
    import os
    fused_only = bool(tables or arrays or reduce or count_flags)
    if fused_only or os.environ.get("DRYAD_EXPR_CODEGEN", "1") != "0":
        from . import codegen
        try:
            return codegen.run_fused(outputs, cols, n, predicate, tables=tables, arrays=arrays,
                                     reduce=reduce, count_flags=count_flags)
        except codegen.CodegenUnavailable as ex:
            if fused_only:
                raise CompileError(f"fused probe/reduce pipeline needs NVRTC code generation: {ex}") from ex
    from ..ops import core as K
    from .ir import retype
    col_dtypes = {k: v.dtype for k, v in cols.items() if isinstance(v, torch.Tensor) and v.dim() == 1}
    kinds = {k: kind_of_dtype(dt) for k, dt in col_dtypes.items()}
    outputs = {k: retype(wrap(e), kinds) for k, e in outputs.items()}
    predicate = retype(predicate, kinds) if predicate is not None else None
    ins, in_cols, out_order, result_reg = compile_program(outputs, predicate, col_dtypes)
    dev = next(iter(cols.values())).device if cols else torch.device("cuda")
    prog = N.ExprProgram()
    prog.ninstr = len(ins)
    prog.ncols_in = len(in_cols)
    prog.ncols_out = len(out_order)
    prog.result_reg = result_reg
    for i, (op, dst, a, b2, imm) in enumerate(ins):
        prog.ins[i].op, prog.ins[i].dst, prog.ins[i].a, prog.ins[i].b, prog.ins[i].imm = op, dst, a, b2, imm
    keep = []
    for i, c in enumerate(in_cols):
        t = cols[c]
        if not t.is_contiguous():
            t = t.contiguous()
        keep.append(t)
        prog.in_[i] = t.data_ptr()
        prog.in_dtype[i] = _DT[t.dtype]
    outs: Dict[str, torch.Tensor] = {}
    for j, (name, kind) in enumerate(out_order):
        dt, code = _OUT_DT[kind]
        t = torch.empty(n, dtype=dt, device=dev)
        outs[name] = t
        prog.out[j] = t.data_ptr()
        prog.out_dtype[j] = code
    flags = torch.empty(n, dtype=torch.uint8, device=dev) if predicate is not None else None
    N.check(N.kernels().dryad_expr_run(prog, n, K.ptr(flags), K.stream()), "expr_run")
    K._count()
    for name, kind in out_order:
        if kind == "b":
            outs[name] = outs[name].view(torch.bool) if outs[name].dtype == torch.uint8 else outs[name]
    return outs, flags
