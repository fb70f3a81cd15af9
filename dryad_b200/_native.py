"""ctypes bindings for the in-tree native libraries.

`kernels()` returns `libdryad_kernels.so` (sm_100a CUDA kernels, C ABI in include/dryad_b200.h);
`host()` returns `libdryad_host.so` (C++ host runtime, C ABI in include/dryad_host.h).
Both are built in-tree by `dryad_b200.build`.  A missing library is an error, never a silent
fallback: GPU ops raise, and only the explicitly CPU code paths (LocalDebug / oracle) run without
the kernel library.
"""
from __future__ import annotations

import ctypes as C
from ctypes import c_int, c_int32, c_int64, c_uint32, c_uint64, c_size_t, c_void_p, c_double
from pathlib import Path

_LIB_DIR = Path(__file__).resolve().parent / "_lib"
_kernels = None
_host = None

MAX_PEERS = 8
FLAG_SLOTS = 8
MAX_LANES = 8
EXPR_MAX_INSTR = 64
EXPR_MAX_COLS = 16


class NativeError(RuntimeError):
    pass


class ShuffleDesc(C.Structure):
    _fields_ = [
        ("ncols", c_int32), ("nparts", c_int32), ("mode", c_int32), ("key_col", c_int32),
        ("key_kind", c_int32), ("key_off", c_int32), ("emit_key", c_int32),
        ("rank", c_int32), ("world", c_int32),
        ("n", c_int64),
        ("src", c_void_p * 8),
        ("width", c_uint32 * 8),
        ("dst_table", c_void_p), ("dst_key_table", c_void_p),
        ("separators", c_void_p), ("capacity", c_void_p),
        ("counts_local", c_void_p), ("counts_peers", c_void_p),
        ("flags_local", c_void_p), ("flags_peers", c_void_p),
        ("epoch", c_uint64),
        ("scratch", c_void_p), ("scratch_bytes", c_size_t),
        ("status", c_void_p),
        ("key_out", c_void_p), ("phases", c_int32), ("barrier_slot", c_int32), ("max_ctas", c_int32),
        ("filter_mod", c_int32), ("filter_rem", c_int32), ("barrier_epoch", c_uint64),
    ]


SHUFFLE_PLAN, SHUFFLE_SCATTER, SHUFFLE_BARRIER = 1, 2, 4


class ExprInstr(C.Structure):
    _fields_ = [("op", c_int32), ("dst", c_int32), ("a", c_int32), ("b", c_int32), ("imm", c_int64)]


class ExprProgram(C.Structure):
    _fields_ = [
        ("ninstr", c_int32), ("ncols_in", c_int32), ("ncols_out", c_int32), ("result_reg", c_int32),
        ("ins", ExprInstr * EXPR_MAX_INSTR),
        ("in_", c_void_p * EXPR_MAX_COLS), ("in_dtype", c_int32 * EXPR_MAX_COLS),
        ("out", c_void_p * EXPR_MAX_COLS), ("out_dtype", c_int32 * EXPR_MAX_COLS),
    ]


def _sig(lib, name, restype, argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = argtypes
    return fn


def _declare_kernels(lib):
    P = c_void_p
    _sig(lib, "dryad_device_count", c_int, [])
    _sig(lib, "dryad_error_string", C.c_char_p, [c_int])
    _sig(lib, "dryad_memcpy_async", c_int, [P, P, c_size_t, c_int, P])
    _sig(lib, "dryad_memset_async", c_int, [P, c_int, c_size_t, P])
    _sig(lib, "dryad_symm_alloc", c_int, [c_size_t, C.POINTER(P)])
    _sig(lib, "dryad_symm_free", c_int, [P])
    _sig(lib, "dryad_ipc_export", c_int, [P, C.c_char_p])
    _sig(lib, "dryad_ipc_open", c_int, [C.c_char_p, C.POINTER(P)])
    _sig(lib, "dryad_ipc_close", c_int, [P])
    _sig(lib, "dryad_signal_barrier", c_int, [P, P, c_int, c_int, c_int, c_uint64, P, P])
    _sig(lib, "dryad_hash_u64", c_int, [P, P, c_int64, P])
    _sig(lib, "dryad_hash_bytes", c_int, [P, P, P, c_int64, P])
    _sig(lib, "dryad_radix_sort_scratch_bytes", c_size_t, [c_int64, c_int, c_int])
    for nm in ("dryad_radix_sort_pairs_u64", "dryad_radix_sort_pairs_u32"):
        _sig(lib, nm, c_int, [P, P, P, P, c_int64, c_int, c_int, c_int, P, c_size_t,
                              C.POINTER(c_int), P])
    _sig(lib, "dryad_shuffle_scratch_bytes", c_size_t, [c_int64, c_int])
    _sig(lib, "dryad_shuffle_run", c_int, [C.POINTER(ShuffleDesc), P])
    _sig(lib, "dryad_sample_keys", c_int, [P, c_int64, c_uint64, c_double, P, P, c_int64, P])
    _sig(lib, "dryad_sample_rows", c_int, [P, c_int, c_int, c_int64, c_uint64, c_double, P, P,
                                           c_int64, P])
    _sig(lib, "dryad_select_separators", c_int, [P, c_int64, c_int, c_int, P, P])
    _sig(lib, "dryad_select_separators_dev", c_int, [P, P, c_int, c_int64, c_int, c_int, P, P])
    _sig(lib, "dryad_extract_be64", c_int, [P, c_int64, c_int, c_int, P, P])
    _sig(lib, "dryad_fix_ties", c_int, [P, P, c_int64, P, c_int, c_int, c_int, c_int, c_int, c_uint64, P, P])
    _sig(lib, "dryad_gather_rows", c_int, [P, P, P, c_int64, c_int, P])
    _sig(lib, "dryad_fix_ties32_keys", c_int, [P, P, c_int64, P, c_int, c_int, P, P, P])
    _sig(lib, "dryad_extract_be_window", c_int, [P, c_int64, c_int, c_int, c_int, P, P])
    _sig(lib, "dryad_shift_u64_to_u32", c_int, [P, c_int64, c_int, P, P])
    _sig(lib, "dryad_fix_ties32", c_int, [P, P, c_int64, P, c_int, c_int, c_int, c_int, c_int, P, P])
    _sig(lib, "dryad_check_sorted_rows", c_int, [P, c_int64, c_int, c_int, c_int, c_int, P, P])
    _sig(lib, "dryad_hash_agg_lane_stride", c_int, [c_int])
    _sig(lib, "dryad_hash_agg_init", c_int, [P, c_uint64, c_int, P, P, P])
    _sig(lib, "dryad_hash_agg_update", c_int, [P, c_uint64, c_int, P, P, P, P, c_int64, c_int, P, P])
    _sig(lib, "dryad_hash_agg_merge_peers", c_int, [P, c_uint64, c_int, P, P, c_int, P, P,
                                                    c_uint64, P, P])
    _sig(lib, "dryad_hash_agg_extract", c_int, [P, c_uint64, c_int, P, P, P, P, P, c_int64, P])
    _sig(lib, "dryad_join_build", c_int, [P, P, c_uint64, P, P, c_int64, P, P])
    _sig(lib, "dryad_join_probe", c_int, [P, P, c_uint64, P, P, c_int64, c_int, P, P, P, P])
    _sig(lib, "dryad_set_contains", c_int, [P, c_uint64, P, c_int64, P, P])
    _sig(lib, "dryad_expr_run", c_int, [C.POINTER(ExprProgram), c_int64, P, P])
    _sig(lib, "dryad_compact_scratch_bytes", c_size_t, [c_int64])
    _sig(lib, "dryad_compact_flags", c_int, [P, c_int64, P, P, P, c_size_t, P])
    _sig(lib, "dryad_exclusive_scan_u32", c_int, [P, c_int64, P, P])
    _sig(lib, "dryad_scan_i64_scratch_bytes", c_size_t, [c_int64])
    _sig(lib, "dryad_exclusive_scan_i64", c_int, [P, c_int64, P, P, c_size_t, P])
    _sig(lib, "dryad_reduce_scratch_bytes", c_size_t, [])
    _sig(lib, "dryad_reduce", c_int, [P, c_int, c_int64, c_int, P, P, c_size_t, P])
    _sig(lib, "dryad_tokenize_scratch_bytes", c_size_t, [c_int64])
    _sig(lib, "dryad_tokenize_count", c_int, [P, c_int64, c_int, c_int, P, P, c_size_t, P])
    _sig(lib, "dryad_tokenize_write", c_int, [P, c_int64, c_int, c_int, P, P, P, P, c_size_t, P])
    _sig(lib, "dryad_hash_views", c_int, [P, P, P, c_int64, P, P])
    _sig(lib, "dryad_str_pack", c_int, [P, P, P, c_int64, c_int, P, P])
    _sig(lib, "dryad_tokenize_verify", c_int, [P, P, P, P, P, c_int64, P, P])
    _sig(lib, "dryad_replicate_slice", c_int, [c_int, c_int, P, P, c_int64, c_int64, P])
    _sig(lib, "dryad_join_table_init", c_int, [P, P, c_uint64, P])
    _sig(lib, "dryad_join_probe_unique", c_int, [P, P, c_uint64, P, c_int64, P, P, P])
    _sig(lib, "dryad_join_build_bcast", c_int, [P, c_int64, c_int64, c_int, P, P, P, c_uint64, c_int, P, P, P, P])
    _sig(lib, "dryad_merge_pairs_u64", c_int, [P, P, c_int64, P, P, c_int64, P, P, P])
    _sig(lib, "dryad_sorted_bounds", c_int, [P, c_int64, P, c_int64, P, P, P])
    _sig(lib, "dryad_emit_ranges", c_int, [P, P, P, c_int64, P, P, P])
    _sig(lib, "dryad_kmeans_norms", c_int, [P, c_int, P, P])
    _sig(lib, "dryad_kmeans_assign", c_int, [P, c_int64, P, P, c_int, P, P, P, P, c_int, P])
    _sig(lib, "dryad_chan_control_bytes", c_size_t, [c_int])
    _sig(lib, "dryad_chan_create", c_int, [P, c_size_t, c_int, c_int, P, C.POINTER(c_void_p)])
    _sig(lib, "dryad_chan_destroy", c_int, [P])
    _sig(lib, "dryad_chan_alloc_block", c_int, [P, c_size_t, c_double, C.POINTER(c_void_p)])
    _sig(lib, "dryad_chan_write_block", c_int, [P, P, c_size_t, P])
    _sig(lib, "dryad_chan_close_writer", c_int, [P])
    _sig(lib, "dryad_chan_read_block", c_int, [P, C.POINTER(c_void_p), C.POINTER(c_size_t), c_double, P])
    _sig(lib, "dryad_chan_release_block", c_int, [P, P, P])
    _sig(lib, "dryad_chan_total_length", c_int64, [P])
    _sig(lib, "dryad_chan_stats", c_int, [P, C.POINTER(c_int64)])
    _sig(lib, "dryad_kmeans_sep", c_int, [P, c_int, P, P])
    _sig(lib, "dryad_kmeans_assign_exact", c_int, [P, c_int64, P, P, P, c_int, P, P, P, P, P, P, P, c_int, P])
    _sig(lib, "dryad_kmeans_update", c_int, [P, P, c_int, c_int, P, P, P, P])
    _sig(lib, "dryad_radix_sort_pairs_u64_noalloc", c_int, [P, P, P, P, c_int64, c_int, c_int, c_int,
                                                            C.POINTER(c_int), P])
    _sig(lib, "dryad_stream_host_func", c_int, [P, P, P])
    _sig(lib, "dryad_ring_push", c_int, [P, c_uint64, P, P, P, P, P, c_uint64, c_int, c_uint64, P, P])
    _sig(lib, "dryad_ring_pop", c_int, [P, c_uint64, P, P, P, P, P, P, c_uint64, c_int, c_uint64, P, P])
    _sig(lib, "dryad_nvls_allreduce_f32", c_int, [P, c_int64, c_int, c_int, P])
    _sig(lib, "dryad_nvls_allreduce_u64", c_int, [P, c_int64, c_int, c_int, P])
    _sig(lib, "dryad_nvls_broadcast", c_int, [P, P, c_int64, P])
    _sig(lib, "dryad_hash_agg_update_auto", c_int, [P, c_uint64, c_int, P, P, P, P, c_int64, c_int, P, P, P])
    _sig(lib, "dryad_hash_agg_update_direct_auto", c_int, [P, c_uint64, c_int, P, c_uint64, P, P, c_int64, P, P, P])
    _sig(lib, "dryad_minmax_i64", c_int, [P, c_int64, P, P])
    _sig(lib, "dryad_hash_agg_direct_lane_stride", c_int, [c_int, P])
    _sig(lib, "dryad_hash_agg_init_direct", c_int, [P, c_uint64, c_int, P, P])
    _sig(lib, "dryad_hash_agg_update_direct", c_int, [P, c_uint64, c_int, P, c_uint64, P, P, c_int64, P, P])
    _sig(lib, "dryad_hash_agg_extract_direct", c_int, [P, c_uint64, c_int, P, c_uint64, P, P, P, c_int64, P])
    _sig(lib, "dryad_hash_agg_merge_peers_direct", c_int, [P, c_uint64, c_int, P, c_uint64, c_int, P, P, c_uint64, P, P])
    _sig(lib, "dryad_hash_agg_update_rows", c_int, [P, c_uint64, c_int, P, P, P, c_int64, c_int, P, P])
    _sig(lib, "dryad_hash_agg_extract_rows", c_int, [P, c_uint64, c_int, P, P, P, P, c_int64, P])
    _sig(lib, "dryad_mark_boundaries", c_int, [P, c_int64, P, P])
    _sig(lib, "dryad_rows_to_columns", c_int, [P, c_int64, c_int, c_int, P, P, P, P])
    _sig(lib, "dryad_columns_to_rows", c_int, [P, c_int64, c_int, c_int, P, P, P, P])
    # optional symbols added by later kernels
    for nm, rt, at in _OPTIONAL:
        if hasattr(lib, nm):
            _sig(lib, nm, rt, at)


_OPTIONAL: list = []


def register_optional(name, restype, argtypes):
    """Let a module declare the signature of a symbol it owns (bound if present)."""
    _OPTIONAL.append((name, restype, argtypes))
    if _kernels is not None and hasattr(_kernels, name):
        _sig(_kernels, name, restype, argtypes)


def kernels_path() -> Path:
    return _LIB_DIR / "libdryad_kernels.so"


def host_path() -> Path:
    """DRYAD_HOST_LIB overrides the library (e.g. an ASan/UBSan or TSan build, bench/sanitize_host.sh)."""
    import os
    alt = os.environ.get("DRYAD_HOST_LIB")
    return Path(alt) if alt else _LIB_DIR / "libdryad_host.so"


def kernels():
    """The CUDA kernel library.  Raises if it has not been built."""
    global _kernels
    if _kernels is None:
        p = kernels_path()
        if not p.exists():
            raise NativeError(
                f"{p} is missing: run `python -m dryad_b200.build` (GPU ops never fall back)")
        lib = C.CDLL(str(p))
        _declare_kernels(lib)
        _kernels = lib
    return _kernels


def host():
    """The C++ host runtime library (built on demand: it only needs g++)."""
    global _host
    if _host is None:
        p = host_path()
        if not p.exists():
            from . import build
            build.build_host()
        lib = C.CDLL(str(p))
        from .runtime import hostlib
        hostlib.declare(lib)
        _host = lib
    return _host


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = kernels().dryad_error_string(rc).decode()
        raise NativeError(f"{what or 'native call'} failed: {msg} (code {rc})")
