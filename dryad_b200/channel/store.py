"""Partitioned tables on disk ("partfile" scheme) — the storage layer.

Reference: PartitionedFileDataProvider (LinqToDryad/DataProvider.cs:378-538), DataPath URI schemes
(LinqToDryad/DataPath.cs:36-124), GM-side partfile streams and commit
(GraphManager/filesystem/DrPartitionFile.cpp:393-402, 561-597).  A table `partfile:///dir/name.pt`
is a text metadata file (base path, partition count, `index,size` lines) plus one data file
`<base>.%08X` per partition.  Data files are block streams written by the native writer
(csrc/host/records.cpp); output is committed atomically: parts are written as `*.tmp`, renamed,
and the `.pt` file is published last.  `file://...txt` sources read plain text lines as LineRecords.
"""
from __future__ import annotations

import ctypes as C
import json
import os
from typing import Any, List, Optional, Sequence

from .. import _native as N
from ..api.errors import DryadLinqException, ErrorCode
from ..api.types import LineRecord
from . import records as R
from .provider import DataProvider, StreamInfo

SCHEMES = ("partfile", "file", "hdfs", "wasb", "azureblob", "hpcdsc")      # the reference's scheme names


class StoredTable:
    def __init__(self, partitions: List[Any], schema):
        self.partitions, self.schema = partitions, schema


def parse_uri(uri: str):
    if "://" not in uri:
        return "partfile", uri
    scheme, rest = uri.split("://", 1)
    DataProvider.get(scheme)               # raises UnrecognizedDataSource for a scheme nobody serves
    path = "/" + rest.lstrip("/") if rest.startswith("/") else rest
    return scheme, path


def _pf_exists(uri: str) -> bool:
    _, path = parse_uri(uri)
    return os.path.exists(path)


def _pf_delete(uri: str) -> None:
    _, path = parse_uri(uri)
    if not os.path.exists(path):
        return
    try:
        base, n, _ = _read_meta(path)
        for i in range(n):
            p = _part_path(base, i)
            if os.path.exists(p):
                os.remove(p)
    except Exception:
        pass
    os.remove(path)
    if os.path.exists(path + ".schema"):
        os.remove(path + ".schema")


def _part_path(base: str, i: int) -> str:
    buf = C.create_string_buffer(4096)
    N.host().dh_partfile_part_path(base.encode(), i, buf, 4096)
    return buf.value.decode()


def _read_meta(path: str):
    lib = N.host()
    base = C.create_string_buffer(4096)
    sizes = (C.c_int64 * 65536)()
    n = lib.dh_partfile_read(path.encode(), base, 4096, sizes, 65536)
    if n < 0:
        raise DryadLinqException(ErrorCode.StreamDoesNotExist, f"cannot read partfile {path} ({n})")
    return base.value.decode(), n, [sizes[i] for i in range(min(n, 65536))]


def write_records_file(path: str, part, compression=None) -> int:
    """One block-stream file of records (the format of table partitions and of file channels between
    vertices).  Returns the stored size; on error the file is removed."""
    lib = N.host()
    comp = compression if isinstance(compression, int) else (1 if str(compression or "None").lower() in ("gzip", "1", "true") else 0)
    try:
        w = lib.dh_writer_open(path.encode(), comp, 0)
        if not w:
            raise DryadLinqException(ErrorCode.FailedToCreateStream, f"cannot create {path}")
        try:
            recs = part.to_records() if hasattr(part, "to_records") else part
            it = iter(recs)
            while True:                               # encode and hand over 64 K records at a time
                import itertools
                chunk = list(itertools.islice(it, 65536))
                if not chunk:
                    break
                data, offs, n = R.encode_many(chunk)
                rc = lib.dh_write_records(w, data, offs, n)
                if rc:
                    raise DryadLinqException(ErrorCode.WriteFileError, f"write failed ({rc})")
        except BaseException:
            lib.dh_writer_close(w, None, None, None)
            raise
        raw, stored, nrec = C.c_int64(), C.c_int64(), C.c_int64()
        rc = lib.dh_writer_close(w, C.byref(raw), C.byref(stored), C.byref(nrec))
        if rc:
            raise DryadLinqException(ErrorCode.WriteFileError, f"close failed ({rc})")
    except BaseException:
        if os.path.exists(path):
            os.remove(path)
        raise
    return stored.value


def _pf_write_part(uri: str, index: int, part, compression: Optional[str] = None):
    """Write ONE partition of a table as an uncommitted version file (SPMD: every rank writes its own
    partition).  Returns (tmp_path, final_path, stored_bytes); nothing is visible until
    `commit_part` renames it and `finish_table` publishes the metadata file
    (GraphManager/filesystem/DrPartitionFile.cpp:561-580, DrGraph::FinalizeGraph)."""
    scheme, path = parse_uri(uri)
    lib = N.host()
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    base = os.path.splitext(path)[0]
    comp = 1 if str(compression or "None").lower() in ("gzip", "1", "true") else 0
    final = _part_path(base, index)
    tmp = final + ".v0.tmp"
    stored = write_records_file(tmp, part, comp)
    return tmp, final, stored


def _pf_commit_part(tmp: str, final: str) -> None:
    if N.host().dh_commit_part(tmp.encode(), final.encode()):
        raise DryadLinqException(ErrorCode.WriteFileError, f"cannot commit {final}")


def _pf_finish_table(uri: str, sizes: Sequence[int], schema=None, compression: Optional[str] = None) -> None:
    """Publish the metadata of a table whose partitions have all been committed."""
    scheme, path = parse_uri(uri)
    base = os.path.splitext(path)[0]
    comp = 1 if str(compression or "None").lower() in ("gzip", "1", "true") else 0
    arr = (C.c_int64 * max(len(sizes), 1))(*sizes)
    rc = N.host().dh_partfile_write(path.encode(), base.encode(), len(sizes), arr, b"")
    if rc:
        raise DryadLinqException(ErrorCode.WriteFileError, f"cannot write {path} ({rc})")
    with open(path + ".schema", "w") as f:
        json.dump({"schema": schema, "compression": "Gzip" if comp else "None"}, f)


def _pf_write_table(uri: str, parts: Sequence[Sequence], schema=None, compression: Optional[str] = None) -> None:
    scheme, path = parse_uri(uri)
    lib = N.host()
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    base = os.path.splitext(path)[0]
    comp = 1 if str(compression or "None").lower() in ("gzip", "1", "true") else 0
    sizes = (C.c_int64 * max(len(parts), 1))()
    tmp_paths = []
    try:
        for i, part in enumerate(parts):
            final = _part_path(base, i)
            tmp = final + ".v0.tmp"
            tmp_paths.append((tmp, final))
            sizes[i] = write_records_file(tmp, part, comp)
        for tmp, final in tmp_paths:                      # commit the successful part versions
            if lib.dh_commit_part(tmp.encode(), final.encode()):
                raise DryadLinqException(ErrorCode.WriteFileError, f"cannot commit {final}")
        rc = lib.dh_partfile_write(path.encode(), base.encode(), len(parts), sizes, b"")
        if rc:
            raise DryadLinqException(ErrorCode.WriteFileError, f"cannot write {path} ({rc})")
        with open(path + ".schema", "w") as f:
            json.dump({"schema": schema, "compression": "Gzip" if comp else "None"}, f)
    except BaseException:
        for tmp, _ in tmp_paths:                          # failed jobs leave no partial outputs
            if os.path.exists(tmp):
                os.remove(tmp)
        raise


def read_partition_file(p: str) -> List:
    lib = N.host()
    r = lib.dh_reader_open(p.encode())
    if not r:
        raise DryadLinqException(ErrorCode.ReadFileError, f"cannot open {p}")
    out: List = []
    ptr = C.POINTER(C.c_uint8)()
    avail = C.c_int64()
    try:
        while True:
            rc = lib.dh_reader_peek(r, C.byref(ptr), C.byref(avail))
            if rc < 0:
                raise DryadLinqException(ErrorCode.FailedToReadFrom, f"corrupt block stream {p} ({rc})")
            if avail.value == 0:
                break
            out.extend(R.decode_many(C.string_at(ptr, avail.value)))   # whole records only (never straddle)
            lib.dh_reader_skip(r, avail.value)
    finally:
        lib.dh_reader_close(r)
    return out


class TextPartition:
    """The lines of (a byte range of) a text file, kept as raw bytes.  The GPU backend uploads the
    bytes as they are (the string column is views into them); LineRecords are only materialised when
    host code iterates (host backend, LocalDebug, enumeration).
    Reference: LinqToDryad/DryadLinqTextReader.cs:36-280."""

    is_text_partition = True                 # cheap type test that does not touch (and so does not load) `data`

    def __init__(self, data, device_buf=None, source=None):
        self._data = data                    # bytes (host copy; b"" when only the device copy exists); None = not read yet
        self.device_buf = device_buf         # optional uint8 tensor already holding the same bytes
        self.source = source                 # (path, lo, hi): the byte range of a file this partition stands for
        self._lines: Optional[List[LineRecord]] = None

    @classmethod
    def from_file(cls, path: str, lo: int, hi: int) -> "TextPartition":
        """A byte range of a text file, read when first used (a vertex that never runs here never reads it)."""
        return cls(None, None, (path, lo, hi))

    @property
    def data(self):
        if self._data is None:
            path, lo, hi = self.source
            with open(path, "rb") as f:
                f.seek(lo)
                self._data = f.read(hi - lo)
        return self._data

    @data.setter
    def data(self, value):
        self._data = value

    def line_count(self) -> int:
        """Number of line terminators, counted once (natively: the vertex statistics ask several times)."""
        if getattr(self, "_nlines", None) is None:
            d = self.data
            self._nlines = int(N.host().hv_count_byte(d, len(d), 10)) if d else 0
        return self._nlines

    @property
    def nbytes(self) -> int:
        if self._data is None:
            return self.source[2] - self.source[1]
        return max(len(self._data), 0 if self.device_buf is None else int(self.device_buf.numel()))

    def to_records(self) -> List[LineRecord]:
        if self._lines is None:
            from ..oracle.text import split_lines
            data = self.data if (self.data or self.device_buf is None) else bytes(self.device_buf.cpu().numpy())
            self._lines = [LineRecord(l.decode("utf-8")) for l in split_lines(data)]
        return self._lines

    def __len__(self):
        return len(self.to_records())

    def __iter__(self):
        return iter(self.to_records())

    def __getitem__(self, i):
        return self.to_records()[i]


class LazyPartitions:
    """The partitions of a stored table, opened on demand: what a job package carries instead of the data.
    Text: (path, byte ranges); partfile: (path of the metadata file).  Indexing loads ONE partition."""

    def __init__(self, uri: str, kind: str, ranges=None, count: int = 0):
        self.uri, self.kind, self.ranges, self.count = uri, kind, ranges, count

    @staticmethod
    def describe(uri: str, partitions):
        try:
            scheme, path = parse_uri(uri)
        except Exception:  # noqa: BLE001
            return None
        if not isinstance(DataProvider.for_uri(uri), PartitionedFileProvider) or not os.path.exists(path):
            return None                                    # other providers: the data travels with the package
        if all(isinstance(p, TextPartition) for p in partitions) and partitions:
            if all(p.source is not None and p.source[0] == path for p in partitions):
                return LazyPartitions(uri, "text", [(p.source[1], p.source[2]) for p in partitions], len(partitions))
            return None
        if path.endswith(".pt") and not any(isinstance(p, TextPartition) for p in partitions):
            return LazyPartitions(uri, "partfile", None, len(partitions))
        return None

    def __len__(self):
        return self.count

    def nbytes(self, i: int) -> int:
        """Stored size of partition i, without reading it."""
        if self.kind == "text":
            return self.ranges[i][1] - self.ranges[i][0]
        _, path = parse_uri(self.uri)
        _, _, sizes = _read_meta(path)
        return int(sizes[i]) if i < len(sizes) else 0

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [self[k] for k in range(*i.indices(self.count))]
        if i < 0:
            i += self.count
        if not 0 <= i < self.count:
            raise IndexError(i)
        _, path = parse_uri(self.uri)
        if self.kind == "text":
            lo, hi = self.ranges[i]
            return TextPartition.from_file(path, lo, hi)
        base, n, _ = _read_meta(path)
        return read_partition_file(_part_path(base, i))

    def __iter__(self):
        return (self[i] for i in range(self.count))


def split_text_file(path: str, n: int) -> List[tuple]:
    """Byte ranges [(lo, hi)] that cut a text FILE into n parts of about equal size at line boundaries,
    reading only a little around each cut."""
    total = os.path.getsize(path)
    if n <= 1 or total == 0:
        return [(0, total)]
    cuts = [0]
    with open(path, "rb") as f:
        for i in range(1, n):
            pos = max(cuts[-1], (total * i) // n)
            f.seek(pos)
            end = total
            while pos < total:
                chunk = f.read(1 << 16)
                if not chunk:
                    break
                k = chunk.find(b"\n")
                if k >= 0:
                    end = pos + k + 1
                    break
                pos += len(chunk)
            cuts.append(min(end, total))
    cuts.append(total)
    return [(cuts[i], cuts[i + 1]) for i in range(n)]


def split_text(data: bytes, n: int) -> List[bytes]:
    """Cut text into n byte ranges of about equal size, each ending at a line boundary."""
    if n <= 1:
        return [data]
    out, pos, total = [], 0, len(data)
    for i in range(n):
        if i == n - 1:
            out.append(data[pos:])
            break
        end = max(pos, min(total, (total * (i + 1)) // n))
        nl = data.find(b"\n", end)
        end = total if nl < 0 else nl + 1
        out.append(data[pos:end])
        pos = end
    return out


def _pf_open_table(uri: str, text_partitions: int = 1) -> StoredTable:
    scheme, path = parse_uri(uri)
    if not os.path.exists(path):
        raise DryadLinqException(ErrorCode.StreamDoesNotExist, f"{uri} does not exist")
    if scheme == "file" or not path.endswith(".pt"):
        # plain text, one LineRecord per line: byte ranges of the file, read when a vertex first touches them
        return StoredTable([TextPartition.from_file(path, lo, hi) for lo, hi in split_text_file(path, text_partitions)],
                           {"value": "s"})
    base, n, _ = _read_meta(path)
    schema = None
    if os.path.exists(path + ".schema"):
        with open(path + ".schema") as f:
            schema = json.load(f).get("schema")
    parts = [read_partition_file(_part_path(base, i)) for i in range(n)]
    return StoredTable(parts, schema)


# ---- providers and the public entry points ---------------------------------------------------------
class PartitionedFileProvider(DataProvider):
    """`partfile://` (and bare paths): metadata file + one block-stream file per partition
    (LinqToDryad/DataProvider.cs:378-538)."""
    scheme = "partfile"
    supports_part_commit = True

    def exists(self, uri):
        return _pf_exists(uri)

    def delete(self, uri):
        _pf_delete(uri)

    def open_table(self, uri, text_partitions=1):
        return _pf_open_table(uri, text_partitions)

    def write_table(self, uri, parts, schema=None, compression=None):
        _pf_write_table(uri, parts, schema, compression)

    def write_part(self, uri, index, part, compression=None):
        return _pf_write_part(uri, index, part, compression)

    def commit_part(self, tmp, final):
        _pf_commit_part(tmp, final)

    def finish_table(self, uri, sizes, schema=None, compression=None):
        _pf_finish_table(uri, sizes, schema, compression)

    def get_stream_info(self, uri):
        _, path = parse_uri(uri)
        if not path.endswith(".pt"):
            return StreamInfo(1, os.path.getsize(path))
        _, n, sizes = _read_meta(path)
        return StreamInfo(n, sum(sizes))

    def get_metadata(self, uri):
        _, path = parse_uri(uri)
        if os.path.exists(path + ".schema"):
            with open(path + ".schema") as f:
                return json.load(f)
        return {"schema": None, "compression": "None"}

    def temporary_uri(self, context, path):
        return f"partfile://{os.path.join(context.work_dir, 'tmp', path)}.pt"


class TextFileProvider(PartitionedFileProvider):
    """`file://`: plain text, one LineRecord per line; writes go to a partfile-format table."""
    scheme = "file"


DataProvider.register(PartitionedFileProvider(), replace=True)
DataProvider.register(TextFileProvider(), replace=True)


def exists(uri: str) -> bool:
    return DataProvider.for_uri(uri).exists(uri)


def delete(uri: str) -> None:
    DataProvider.for_uri(uri).delete(uri)


def open_table(uri: str, text_partitions: int = 1):
    return DataProvider.for_uri(uri).open_table(uri, text_partitions)


def write_table(uri: str, parts: Sequence[Sequence], schema=None, compression: Optional[str] = None) -> None:
    DataProvider.for_uri(uri).write_table(uri, parts, schema, compression)


def supports_part_commit(uri: str) -> bool:
    return DataProvider.for_uri(uri).supports_part_commit


def write_part(uri: str, index: int, part, compression: Optional[str] = None):
    return DataProvider.for_uri(uri).write_part(uri, index, part, compression)


def commit_part(tmp: str, final: str, uri: Optional[str] = None) -> None:
    (DataProvider.for_uri(uri) if uri else DataProvider.get("partfile")).commit_part(tmp, final)


def finish_table(uri: str, sizes: Sequence[int], schema=None, compression: Optional[str] = None) -> None:
    DataProvider.for_uri(uri).finish_table(uri, sizes, schema, compression)


def stream_info(uri: str) -> StreamInfo:
    return DataProvider.for_uri(uri).get_stream_info(uri)

