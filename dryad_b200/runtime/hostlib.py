"""ctypes signatures of libdryad_host.so (include/dryad_host.h)."""
from __future__ import annotations

import ctypes as C
from ctypes import c_char_p, c_double, c_int, c_int32, c_int64, c_uint8, c_void_p

_DECLS = []


def _d(name, restype, argtypes):
    _DECLS.append((name, restype, argtypes))


P = c_void_p
_d("dh_writer_open", P, [c_char_p, c_int, c_int64])
_d("dh_write_bytes", None, [P, c_char_p, c_int64])
_d("dh_write_u8", None, [P, c_uint8])
_d("dh_write_i32", None, [P, c_int32])
_d("dh_write_i64", None, [P, c_int64])
_d("dh_write_f64", None, [P, c_double])
_d("dh_write_compact", None, [P, c_int32])
_d("dh_write_string", None, [P, c_char_p, c_int32, c_int32])
_d("dh_writer_end_record", c_int, [P])
_d("dh_write_records", c_int, [P, c_char_p, c_char_p, c_int64])
_d("dh_writer_write_rows", c_int, [P, P, c_int64, c_int32])
_d("dh_writer_close", c_int, [P, C.POINTER(c_int64), C.POINTER(c_int64), C.POINTER(c_int64)])
_d("dh_reader_open", P, [c_char_p])
_d("dh_reader_at_end", c_int, [P])
_d("dh_read_bytes", c_int, [P, P, c_int64])
_d("dh_read_u8", c_int, [P, C.POINTER(c_uint8)])
_d("dh_read_i32", c_int, [P, C.POINTER(c_int32)])
_d("dh_read_i64", c_int, [P, C.POINTER(c_int64)])
_d("dh_read_f64", c_int, [P, C.POINTER(c_double)])
_d("dh_read_compact", c_int, [P, C.POINTER(c_int32)])
_d("dh_reader_peek", c_int, [P, C.POINTER(C.POINTER(c_uint8)), C.POINTER(c_int64)])
_d("dh_reader_skip", c_int, [P, c_int64])
_d("dh_reader_close", c_int, [P])
_d("dh_gzip_bound", c_int64, [c_int64])
_d("dh_gzip_compress", c_int64, [P, c_int64, P, c_int64])
_d("dh_gzip_decompress", c_int64, [P, c_int64, P, c_int64])
_d("dh_partfile_write", c_int, [c_char_p, c_char_p, c_int32, C.POINTER(c_int64), c_char_p])
_d("dh_partfile_read", c_int, [c_char_p, c_char_p, c_int32, C.POINTER(c_int64), c_int32])
_d("dh_partfile_part_path", None, [c_char_p, c_int32, c_char_p, c_int32])
_d("dh_commit_part", c_int, [c_char_p, c_char_p])


def declare(lib) -> None:
    for name, rt, at in _DECLS:
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = rt, at


def add(name, restype, argtypes):
    """Modules owning further host symbols (scheduler, mailbox, …) register them here."""
    _d(name, restype, argtypes)

# ---- graph manager / dynamic managers -----------------------------------------------------------
from ctypes import c_uint64  # noqa: E402

_d("gm_create", P, [c_int, c_int])
_d("gm_destroy", None, [P])
_d("gm_set_outlier_params", None, [P, c_double, c_double, c_double])
_d("gm_add_stage", c_int, [P, c_char_p, c_int])
_d("gm_connect", c_int, [P, c_int, c_int, c_int])
_d("gm_connect_ex", c_int, [P, c_int, c_int, c_int, c_int])
_d("gm_resize_stage", c_int, [P, c_int, c_int, c_int])
_d("gm_splice", c_int, [P, c_int, C.POINTER(c_int), c_int, C.POINTER(c_int), c_int])
_d("gm_vertex_inputs", c_int, [P, c_int, C.POINTER(c_int), c_int])
_d("gm_stage_vertices", c_int, [P, c_int, C.POINTER(c_int), c_int])
_d("gm_vertex_clique", c_int, [P, c_int, C.POINTER(c_int), c_int])
_d("gm_live_vertex_count", c_int, [P])
_d("gm_gang_restarts", c_int, [P])
_d("gm_start", c_int, [P])
_d("gm_next_ready", c_int, [P, c_double, C.POINTER(c_int), C.POINTER(c_int)])
_d("gm_report_completed", c_int, [P, c_int, c_int, c_double, c_double])
_d("gm_report_failed", c_int, [P, c_int, c_int, c_int, c_char_p])
_d("gm_check_duplicates", c_int, [P, c_double])
_d("gm_job_state", c_int, [P])
_d("gm_vertex_count", c_int, [P])
_d("gm_vertex_info", c_int, [P, c_int] + [C.POINTER(c_int)] * 5)
_d("gm_counters", c_int, [P] + [C.POINTER(c_int)] * 4)
_d("gm_error", c_int, [P, c_char_p, c_int])
_d("gm_stage_model", c_int, [P, c_int] + [C.POINTER(c_double)] * 4 + [c_double])
_d("dyn_aggregate_groups", c_int, [c_int, C.POINTER(c_int64), C.POINTER(c_int32), C.POINTER(c_int32), c_int, c_int64,
                                   c_int64, C.POINTER(c_int32)])
_d("dyn_consumer_count", c_int, [c_int64, c_int64, c_int, c_int])
_d("dyn_broadcast_tree", c_int, [c_int, C.POINTER(c_int32)])
_d("dyn_range_consumer_count", c_int, [c_int64, c_double, c_int64, c_int, c_int])
PUMP_HANDLER = C.CFUNCTYPE(None, c_int64, c_int64, P)
_d("pump_create", P, [c_int, PUMP_HANDLER, P])
_d("pump_post", None, [P, c_int64, c_int64, c_double])
_d("pump_wait_delivered", c_int64, [P, c_int64, c_double])
_d("pump_make_notifier", P, [P, c_int64, c_int64])
_d("pump_notifier_fn", P, [])
_d("pump_destroy", None, [P])
_d("mbox_create", P, [])
_d("mbox_destroy", None, [P])
_d("mbox_set", c_uint64, [P, c_char_p, c_char_p, c_char_p, c_int64])
_d("mbox_block_on_status", c_uint64, [P, c_char_p, c_uint64, c_double, c_char_p, c_int32, c_char_p, c_int64,
                                      C.POINTER(c_int64)])
_d("ls_create", P, [c_int, C.POINTER(c_int32), c_double, c_double])
_d("ls_destroy", None, [P])
_d("ls_submit", None, [P, c_int64, C.POINTER(c_int32), c_int, c_int])
_d("ls_computer_idle", c_int64, [P, c_int, c_double])
_d("ls_waiting", c_int, [P])
_d("fifo_create", P, [c_int])
_d("fifo_destroy", None, [P])
_d("fifo_write", c_int, [P, c_char_p, c_int64, c_double])
_d("fifo_close_writer", None, [P])
_d("fifo_abort", None, [P])
_d("fifo_read", c_int64, [P, c_char_p, c_int64, c_double])
_d("fifo_depth_now", c_int, [P])
_d("cache_create", P, [c_char_p, c_int64, c_int64])
_d("cache_destroy", None, [P])
_d("cache_put", c_int, [P, c_char_p, c_char_p, c_int64])
_d("cache_get", c_int64, [P, c_char_p, c_char_p, c_int64])
_d("cache_stats", c_int, [P] + [C.POINTER(c_int64)] * 4)
_d("fprint64", c_uint64, [c_uint64, c_char_p, c_int64])
_d("hv_split_count", P, [c_char_p, c_int64, c_int, c_int, c_int])
_d("hv_count_byte", c_int64, [c_char_p, c_int64, c_int])
_d("hv_groups", c_int64, [P])
_d("hv_word_bytes", c_int64, [P])
_d("hv_total", c_int64, [P])
_d("hv_export", None, [P, c_char_p, C.POINTER(c_int64), C.POINTER(c_int64)])
_d("hv_free", None, [P])
_d("cpu_wordcount", c_int64, [c_char_p, c_int64, c_int, c_int, C.POINTER(c_int64)])
_d("cpu_sort_rows", c_int, [P, c_int64, c_int, c_int, c_int, P])
_d("cpu_groupby_sum_count", c_int64, [P, P, c_int64, c_int, C.POINTER(c_int64)])
_d("vc_create", P, [c_int, c_int, c_int])
_d("vc_destroy", None, [P])
_d("vc_add_arg", None, [P, c_char_p])
_d("vc_add_input", None, [P, c_char_p, c_int64])
_d("vc_add_output", None, [P, c_char_p, c_int64])
_d("vc_set_status", None, [P, c_int, c_int, c_char_p])
_d("vc_add_channel_status", None, [P, c_int64, c_int])
_d("vc_serialize", c_int64, [P, c_char_p, c_int64])
_d("vc_parse", P, [c_char_p, c_int64])
_d("vc_get_header", c_int, [P] + [C.POINTER(c_int)] * 5)
_d("vc_count", c_int, [P, c_int])
_d("vc_get_item", c_int64, [P, c_int, c_int, c_char_p, c_int])
_d("affinity_merge", c_int, [c_int, C.POINTER(c_int32), C.POINTER(c_int64), C.POINTER(c_int32), c_int, c_double, C.POINTER(c_int32)])
