"""Record (de)serialisation on the reference's wire primitives.

Primitives (byte-exact with csrc/host/records.cpp and the reference
LinqToDryad/DryadLinqBinaryWriter.cs:336-546): little-endian fixed-width numbers; compact int =
1 byte below 0x80 else 4 bytes big-endian with the top bit set; string = compact(char count) +
compact(UTF-8 byte count) + bytes.

The reference generates one serializer class per record type (DryadLinqCodeGen.cs:496-1479).  The
Python analog is a tagged self-describing encoding for arbitrary records plus a fixed-width fast
path for numeric columnar schemas (rows of packed little-endian fields).
"""
from __future__ import annotations

import struct
from typing import Any, List, Tuple

from ..api.types import ForkTuple, Grouping, LineRecord, Pair, Record

T_NONE, T_FALSE, T_TRUE, T_INT, T_FLOAT, T_STR, T_BYTES, T_TUPLE, T_LIST, T_RECORD, T_PAIR, T_GROUP, T_LINE, T_BIGINT, T_DICT, T_FORK = range(16)
T_GUID, T_DATETIME, T_DECIMAL, T_FORKVALUE, T_CUSTOM = 16, 17, 18, 19, 20
T_ENUM, T_OBJECT = 21, 22

# ---- automatic serialization of user types (the reference generates a serializer per record type from
#      its public fields: DryadLinqCodeGen.cs:496-1479; rules and errors DryadLinqSerialization.cs,
#      SerializationTests.cs: nesting, arrays, enums; circular types are an error) -----------------------
_encoding_now: set = set()
_PUBLIC: dict = {}


def _type_name(cls) -> str:
    return f"{cls.__module__}:{cls.__qualname__}"


def _resolve_type(name: str):
    import importlib
    from ..api.errors import DryadLinqException, ErrorCode
    mod, _, qual = name.partition(":")
    if "<locals>" in qual:
        raise DryadLinqException(ErrorCode.TypeRequiredToBePublic,
                                 f"records of type {name} cannot be read back: the class is defined inside a function")
    try:
        obj = importlib.import_module(mod)
        for part in qual.split("."):
            obj = getattr(obj, part)
        return obj
    except Exception as ex:  # noqa: BLE001
        raise DryadLinqException(ErrorCode.FailedToDeserialize, f"cannot find type {name}: {ex}") from ex


def _fields_of(obj):
    """Field name -> value of a user object: dataclass fields, else __dict__, else __slots__."""
    import dataclasses
    if dataclasses.is_dataclass(obj):
        return {f.name: getattr(obj, f.name) for f in dataclasses.fields(obj)}
    d = getattr(obj, "__dict__", None)
    if d is not None:
        return dict(d)
    names = [n for c in type(obj).__mro__ for n in getattr(c, "__slots__", ())]
    if names:
        return {n: getattr(obj, n) for n in names if hasattr(obj, n)}
    return None


def _encode_object(obj, out: List[bytes]) -> bool:
    import enum
    import types
    from ..api.errors import DryadLinqException, ErrorCode
    cls = type(obj)
    if isinstance(obj, enum.Enum):
        out.append(bytes((T_ENUM,)) + enc_string(_type_name(cls)))
        encode(obj.value, out)
        return True
    if isinstance(obj, (types.FunctionType, types.BuiltinFunctionType, types.MethodType, types.ModuleType, type)) \
            or cls.__module__ == "builtins":
        return False
    fields = _fields_of(obj)
    if fields is None:
        return False
    ok = _PUBLIC.get(cls)
    if ok is None:
        # another process must find the class again by name (the reference requires record types to be public):
        # it has to be reachable as module.qualname, and that name has to lead back to this very class
        ok = False
        if "<locals>" not in cls.__qualname__ and cls.__module__ != "__main__":
            try:
                ok = _resolve_type(_type_name(cls)) is cls
            except DryadLinqException:
                ok = False
        _PUBLIC[cls] = ok
    if not ok:
        raise DryadLinqException(ErrorCode.TypeRequiredToBePublic,
                                 f"records of type {cls.__qualname__} cannot be serialised: define the class at module level "
                                 "of an importable module")
    if id(obj) in _encoding_now:
        raise DryadLinqException(ErrorCode.CannotHandleCircularTypes,
                                 f"records of type {cls.__name__} refer to themselves: circular types cannot be serialised")
    _encoding_now.add(id(obj))
    try:
        out.append(bytes((T_OBJECT,)) + enc_string(_type_name(cls)) + compact(len(fields)))
        for k, v in fields.items():
            out.append(enc_string(k))
            encode(v, out)
    finally:
        _encoding_now.discard(id(obj))
    return True

# ---- custom serializers ([CustomDryadLinqSerializer], LinqToDryad/Attributes.cs:37-275;
#      IDryadLinqSerializer<T>.Read/Write, DryadLinqSerialization.cs:41-789) --------------------------
_CUSTOM_BY_TYPE: dict = {}
_CUSTOM_BY_NAME: dict = {}


def register_serializer(cls, name: str, write, read) -> None:
    """`write(obj) -> bytes`, `read(bytes) -> obj`.  `name` travels in the stream, so the reading
    process must have registered the same name (the reference ships the generated serializer
    assembly with the job; here the job's modules are imported on every worker)."""
    _CUSTOM_BY_TYPE[cls] = (name, write)
    _CUSTOM_BY_NAME[name] = read


def custom_serializer(write, read, name: str = ""):
    """Class decorator form: `@custom_serializer(write=..., read=...)`."""
    def mark(cls):
        register_serializer(cls, name or f"{cls.__module__}.{cls.__qualname__}", write, read)
        return cls
    return mark


def compact(v: int) -> bytes:
    if 0 <= v < 0x80:
        return bytes((v,))
    if not 0 <= v < (1 << 31):
        raise ValueError("compact int out of range")
    return bytes(((v >> 24) | 0x80, (v >> 16) & 0xFF, (v >> 8) & 0xFF, v & 0xFF))


def read_compact(buf: memoryview, pos: int) -> Tuple[int, int]:
    b0 = buf[pos]
    if b0 < 0x80:
        return b0, pos + 1
    return ((b0 & 0x7F) << 24) | (buf[pos + 1] << 16) | (buf[pos + 2] << 8) | buf[pos + 3], pos + 4


def enc_string(s: str) -> bytes:
    b = s.encode("utf-8")
    return compact(len(s)) + compact(len(b)) + b


def encode(obj: Any, out: List[bytes]) -> None:
    if obj is None:
        out.append(bytes((T_NONE,)))
    elif obj is True:
        out.append(bytes((T_TRUE,)))
    elif obj is False:
        out.append(bytes((T_FALSE,)))
    elif isinstance(obj, int):
        if -(1 << 63) <= obj < (1 << 63):
            out.append(bytes((T_INT,)) + struct.pack("<q", obj))
        else:
            s = str(obj)
            out.append(bytes((T_BIGINT,)) + enc_string(s))
    elif isinstance(obj, float):
        out.append(bytes((T_FLOAT,)) + struct.pack("<d", obj))
    elif isinstance(obj, LineRecord):
        out.append(bytes((T_LINE,)) + enc_string(str(obj)))
    elif isinstance(obj, str):
        out.append(bytes((T_STR,)) + enc_string(obj))
    elif isinstance(obj, (bytes, bytearray)):
        out.append(bytes((T_BYTES,)) + compact(len(obj)) + bytes(obj))
    elif isinstance(obj, Grouping):
        out.append(bytes((T_GROUP,)))
        encode(obj.key, out)
        out.append(compact(len(obj)))
        for x in obj:
            encode(x, out)
    elif isinstance(obj, Pair):
        out.append(bytes((T_PAIR,)))
        encode(obj[0], out)
        encode(obj[1], out)
    elif isinstance(obj, ForkTuple):
        out.append(bytes((T_FORK,)) + compact(len(obj)))
        for x in obj:
            encode(x, out)
    elif isinstance(obj, Record):
        out.append(bytes((T_RECORD,)) + compact(len(obj)))
        for k, v in obj.items():
            out.append(enc_string(k))
            encode(v, out)
    elif isinstance(obj, dict):
        out.append(bytes((T_DICT,)) + compact(len(obj)))
        for k, v in obj.items():
            encode(k, out)
            encode(v, out)
    elif isinstance(obj, tuple):
        out.append(bytes((T_TUPLE,)) + compact(len(obj)))
        for x in obj:
            encode(x, out)
    elif isinstance(obj, list):
        out.append(bytes((T_LIST,)) + compact(len(obj)))
        for x in obj:
            encode(x, out)
    elif type(obj) in _CUSTOM_BY_TYPE:
        name, write = _CUSTOM_BY_TYPE[type(obj)]
        body = write(obj)
        out.append(bytes((T_CUSTOM,)) + enc_string(name) + compact(len(body)) + bytes(body))
    elif _is_builtin_extra(obj):
        _encode_extra(obj, out)
    else:
        import numpy as np
        if isinstance(obj, np.integer):
            encode(int(obj), out)
        elif isinstance(obj, np.floating):
            encode(float(obj), out)
        elif isinstance(obj, np.bool_):
            encode(bool(obj), out)
        elif _encode_object(obj, out):
            pass
        else:
            from ..api.errors import DryadLinqException, ErrorCode
            raise DryadLinqException(ErrorCode.TypeNotSerializable,
                                     f"records of type {type(obj).__name__} cannot be serialised")


def _is_builtin_extra(obj) -> bool:
    import datetime
    import decimal
    import uuid
    from ..api.types import ForkValue
    return isinstance(obj, (uuid.UUID, datetime.datetime, decimal.Decimal, ForkValue))


def _encode_extra(obj, out: List[bytes]) -> None:
    """Guid (16 bytes, .NET field order = bytes_le), DateTime (int64 ticks of 100 ns since
    0001-01-01, the reference's SqlDateTime/DateTime writers), Decimal (string form), ForkValue."""
    import datetime
    import decimal
    import uuid
    from ..api.types import ForkValue
    if isinstance(obj, uuid.UUID):
        out.append(bytes((T_GUID,)) + obj.bytes_le)
    elif isinstance(obj, datetime.datetime):
        naive = obj.replace(tzinfo=None) - (obj.utcoffset() or datetime.timedelta(0))
        delta = naive - datetime.datetime(1, 1, 1)
        ticks = (delta.days * 86400 + delta.seconds) * 10_000_000 + delta.microseconds * 10
        out.append(bytes((T_DATETIME,)) + struct.pack("<q", ticks))
    elif isinstance(obj, decimal.Decimal):
        out.append(bytes((T_DECIMAL,)) + enc_string(str(obj)))
    elif isinstance(obj, ForkValue):
        out.append(bytes((T_FORKVALUE, 1 if obj.has_value else 0)))
        if obj.has_value:
            encode(obj.value, out)


def _encode_record_py(obj: Any) -> bytes:
    out: List[bytes] = []
    encode(obj, out)
    return b"".join(out)


def _load_native():
    """The CPython extension csrc/py/pyrecords.cpp (built into dryad_b200/_lib by `python -m dryad_b200.build`):
    the same wire format at C speed for the common record shapes; everything else comes back to the
    functions of this module value by value.  Absent (no Python headers at build time): pure Python."""
    import importlib.util
    import os
    if os.environ.get("DRYAD_NATIVE_RECORDS", "1") == "0":
        return None
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_lib", "_dryad_records.so")
    if not os.path.exists(path):
        return None
    try:
        spec = importlib.util.spec_from_file_location("_dryad_records", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.configure(Record, Pair, LineRecord, _encode_record_py, lambda buf, pos: decode(memoryview(buf), pos))
        return mod
    except Exception:  # noqa: BLE001  (an extension built for another interpreter: fall back)
        return None


def encode_record(obj: Any) -> bytes:
    return _native.encode_one(obj) if _native is not None else _encode_record_py(obj)


def encode_many(records) -> Tuple[bytes, bytes, int]:
    """-> (data, offsets, n): the records back to back, `offsets` = int64[n + 1] as bytes."""
    if _native is not None:
        data, offs = _native.encode_many(records)
        return data, offs, len(offs) // 8 - 1
    parts = [_encode_record_py(r) for r in records]
    pos, offs = 0, [0]
    for b in parts:
        pos += len(b)
        offs.append(pos)
    return b"".join(parts), struct.pack(f"<{len(offs)}q", *offs), len(parts)


def decode_many(buf) -> List[Any]:
    """Every record of a buffer that holds whole records back to back."""
    if _native is not None:
        return _native.decode_many(buf)
    mv = memoryview(buf)
    out, pos, n = [], 0, len(mv)
    while pos < n:
        obj, pos = decode(mv, pos)
        out.append(obj)
    return out


def _dec_string(buf: memoryview, pos: int) -> Tuple[str, int]:
    _, pos = read_compact(buf, pos)
    nb, pos = read_compact(buf, pos)
    return bytes(buf[pos:pos + nb]).decode("utf-8"), pos + nb


def decode(buf: memoryview, pos: int) -> Tuple[Any, int]:
    t = buf[pos]
    pos += 1
    if t == T_NONE:
        return None, pos
    if t == T_TRUE:
        return True, pos
    if t == T_FALSE:
        return False, pos
    if t == T_INT:
        return struct.unpack_from("<q", buf, pos)[0], pos + 8
    if t == T_FLOAT:
        return struct.unpack_from("<d", buf, pos)[0], pos + 8
    if t == T_STR:
        return _dec_string(buf, pos)
    if t == T_LINE:
        s, pos = _dec_string(buf, pos)
        return LineRecord(s), pos
    if t == T_BIGINT:
        s, pos = _dec_string(buf, pos)
        return int(s), pos
    if t == T_BYTES:
        n, pos = read_compact(buf, pos)
        return bytes(buf[pos:pos + n]), pos + n
    if t in (T_TUPLE, T_LIST, T_FORK):
        n, pos = read_compact(buf, pos)
        items = []
        for _ in range(n):
            v, pos = decode(buf, pos)
            items.append(v)
        return (tuple(items) if t == T_TUPLE else ForkTuple(*items) if t == T_FORK else items), pos
    if t == T_PAIR:
        a, pos = decode(buf, pos)
        b, pos = decode(buf, pos)
        return Pair(a, b), pos
    if t == T_GROUP:
        k, pos = decode(buf, pos)
        n, pos = read_compact(buf, pos)
        g = Grouping(k)
        for _ in range(n):
            v, pos = decode(buf, pos)
            g.append(v)
        return g, pos
    if t == T_RECORD:
        n, pos = read_compact(buf, pos)
        r = Record()
        for _ in range(n):
            k, pos = _dec_string(buf, pos)
            v, pos = decode(buf, pos)
            r[k] = v
        return r, pos
    if t == T_DICT:
        n, pos = read_compact(buf, pos)
        d = {}
        for _ in range(n):
            k, pos = decode(buf, pos)
            v, pos = decode(buf, pos)
            d[k] = v
        return d, pos
    if t == T_GUID:
        import uuid
        return uuid.UUID(bytes_le=bytes(buf[pos:pos + 16])), pos + 16
    if t == T_DATETIME:
        import datetime
        ticks = struct.unpack_from("<q", buf, pos)[0]
        return datetime.datetime(1, 1, 1) + datetime.timedelta(microseconds=ticks // 10), pos + 8
    if t == T_DECIMAL:
        import decimal
        sv, pos = _dec_string(buf, pos)
        return decimal.Decimal(sv), pos
    if t == T_FORKVALUE:
        from ..api.types import ForkValue
        has = buf[pos]
        pos += 1
        if not has:
            return ForkValue(), pos
        v, pos = decode(buf, pos)
        return ForkValue(v), pos
    if t == T_ENUM:
        name, pos = _dec_string(buf, pos)
        v, pos = decode(buf, pos)
        return _resolve_type(name)(v), pos
    if t == T_OBJECT:
        name, pos = _dec_string(buf, pos)
        n, pos = read_compact(buf, pos)
        cls = _resolve_type(name)
        obj = cls.__new__(cls)
        for _ in range(n):
            k, pos = _dec_string(buf, pos)
            v, pos = decode(buf, pos)
            object.__setattr__(obj, k, v)
        return obj, pos
    if t == T_CUSTOM:
        name, pos = _dec_string(buf, pos)
        n, pos = read_compact(buf, pos)
        read = _CUSTOM_BY_NAME.get(name)
        from ..api.errors import DryadLinqException, ErrorCode
        if read is None:
            raise DryadLinqException(ErrorCode.FailedToDeserialize, f"no serializer registered for {name!r}")
        return read(bytes(buf[pos:pos + n])), pos + n
    from ..api.errors import DryadLinqException, ErrorCode
    raise DryadLinqException(ErrorCode.FailedToDeserialize, f"unknown record tag {t}")


_native = _load_native()

