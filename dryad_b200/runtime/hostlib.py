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
