"""Column-expression IR + tracing of user lambdas (see ir.py) and GPU lowering (compile.py)."""
from .ir import Expr, RowProxy, TraceError, cast, col, eval_numpy, eval_row, gather, lit, probe, probe_slot, retype, rowid, to_string, trace, where, wrap

__all__ = ["Expr", "RowProxy", "TraceError", "cast", "col", "eval_numpy", "eval_row", "gather", "lit", "probe", "probe_slot", "retype", "rowid",
           "to_string", "trace", "where", "wrap"]
