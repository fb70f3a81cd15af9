"""Channel endpoints by URI scheme (reference: the scheme factory of
DryadVertex/VertexHost/system/channel/src/concreterchannel.cpp:45-50, 1147-1230, 1334-1370 —
file://, fifo://, null:// plus managed schemes; naming of intermediate channels
`<vertexId>_<port>_<version>.tmp`, GraphManager/vertex/DrOutputGenerator.cpp:221-241).

Writers and readers move BLOCKS of serialised records (whole records per block).  `hbm://` names
a partition buffer that stays in device memory (the default channel between GPU stages);
`file://` is a native block stream on disk; `fifo://<depth>/<name>` is a bounded in-process FIFO
(depth 32 = blocking, 4294967295 = unbounded); `null://` discards.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Dict, Iterator, List, Optional

from .. import _native as N
from . import records as R
from .fifo import Fifo

_fifos: Dict[str, Fifo] = {}
_hbm: Dict[str, object] = {}
_lock = threading.Lock()


def channel_name(vertex_id: int, port: int, version: int) -> str:
    return f"{vertex_id}_{port}_{version}.tmp"


def fifo_uri(vertex_id: int, port: int, version: int, blocking: bool = True) -> str:
    return f"fifo://{32 if blocking else 4294967295}/{vertex_id}_{port}_{version}"


class ChannelWriter:
    def __init__(self, uri: str, compression=None, timeout: float = 60.0):
        """`compression`: gzip the blocks of a file channel (the intermediate-data compression setting);
        `timeout`: how long a FIFO writer waits for the reader to make room."""
        self.uri = uri
        self.timeout = timeout
        self.scheme, _, rest = uri.partition("://")
        self.records = 0
        self.bytes = 0
        self._buf: List[bytes] = []
        self._w = None
        if self.scheme == "file":
            comp = compression if isinstance(compression, int) else (1 if str(compression or "None").lower() in ("gzip", "1", "true") else 0)
            self._w = N.host().dh_writer_open(("/" + rest.lstrip("/")).encode(), comp, 0)
            if not self._w:
                raise OSError(f"cannot create {uri}")
        elif self.scheme == "fifo":
            depth, _, name = rest.partition("/")
            with _lock:
                self._fifo = _fifos.setdefault(name, Fifo(0 if int(depth) >= 4294967295 else int(depth)))
        elif self.scheme in ("null", "hbm"):
            pass
        else:
            raise ValueError(f"unknown channel scheme {self.scheme!r}")

    def write(self, record) -> None:
        b = R.encode_record(record)
        self.records += 1
        self.bytes += len(b)
        if self.scheme == "file":
            lib = N.host()
            lib.dh_write_bytes(self._w, b, len(b))
            lib.dh_writer_end_record(self._w)
        elif self.scheme == "fifo":
            self._buf.append(b)
            if sum(map(len, self._buf)) >= 64 * 1024:
                self._flush_fifo()
        elif self.scheme == "hbm":
            self._buf.append(record)

    def _flush_fifo(self):
        if self._buf:
            self._fifo.write(b"".join(self._buf), timeout=self.timeout)
            self._buf = []

    def close(self) -> None:
        if self.scheme == "file":
            if self._w:
                stored = C.c_int64()
                rc = N.host().dh_writer_close(self._w, None, C.byref(stored), None)
                self._w = None
                self.stored_bytes = stored.value
                if rc:
                    raise OSError(f"closing {self.uri} failed ({rc})")
        elif self.scheme == "fifo":
            self._flush_fifo()
            self._fifo.close_writer()
        elif self.scheme == "hbm":
            with _lock:
                _hbm[self.uri] = self._buf


class ChannelReader:
    def __init__(self, uri: str, timeout: float = 60.0):
        self.uri = uri
        self.timeout = timeout
        self.scheme, _, rest = uri.partition("://")
        self._rest = rest

    def __iter__(self) -> Iterator:
        if self.scheme == "file":
            from .store import read_partition_file
            yield from read_partition_file("/" + self._rest.lstrip("/"))
        elif self.scheme == "fifo":
            _, _, name = self._rest.partition("/")
            with _lock:
                f = _fifos.setdefault(name, Fifo(32))
            while True:
                blk = f.read(timeout=self.timeout)
                if blk is None:
                    break
                buf = memoryview(blk)
                pos = 0
                while pos < len(buf):
                    obj, pos = R.decode(buf, pos)
                    yield obj
        elif self.scheme == "hbm":
            with _lock:
                data = _hbm.get(self.uri, [])
            yield from data
        elif self.scheme == "null":
            return
        else:
            raise ValueError(f"unknown channel scheme {self.scheme!r}")


class VertexReader:
    """Multi-port input of a vertex (DryadLinqVertexReader, LinqToDryad/DryadLinqVertexReader.cs:41-257):
    one record reader per input channel, read one port after the other.  Unless `keep_port_order` the
    ports are visited in a random permutation, so that the consumers of a CrossProduct edge do not all
    pull from producer 0 first (`VertexReader.cs:67-83`); the permutation is seeded by the vertex id,
    which makes a retried vertex read in the same order."""

    def __init__(self, uris: List[str], keep_port_order: bool = True, seed: int = 0):
        self.ports = [ChannelReader(u) for u in uris]
        self.order = list(range(len(self.ports)))
        if not keep_port_order and len(self.order) > 1:
            import random
            random.Random(seed).shuffle(self.order)

    def port(self, i: int) -> ChannelReader:
        return self.ports[i]

    def __iter__(self) -> Iterator:
        for i in self.order:
            yield from self.ports[i]


class VertexWriter:
    """Multi-port output of a vertex (DryadLinqVertexWriter, LinqToDryad/DryadLinqVertexWriter.cs:41-187):
    distributors write each record to the port its partition function picked (`WriteItem(item, port)`),
    Fork hands one record writer per branch (`GetRecordWriter(port)`)."""

    def __init__(self, uris: List[str], compression=None):
        self.ports = [ChannelWriter(u, compression) for u in uris]

    def write_item(self, item, port: int) -> None:
        self.ports[port].write(item)

    def write_item_sequence(self, items, port: int = 0) -> None:
        w = self.ports[port]
        for it in items:
            w.write(it)

    def get_record_writer(self, port: int) -> ChannelWriter:
        return self.ports[port]

    @property
    def records(self) -> List[int]:
        return [w.records for w in self.ports]

    def close(self) -> None:
        for w in self.ports:
            w.close()


def ensure_fifo(name: str, depth: int = 32) -> None:
    """Create a named FIFO before either end touches it, so that an abort by one end is seen by the other
    even when it comes first."""
    with _lock:
        if name not in _fifos:
            _fifos[name] = Fifo(0 if depth >= 4294967295 else depth)


def drop_fifo(name: str, abort: bool = False) -> None:
    """Forget a named FIFO once both ends are done with it (or abort it: both ends fail fast)."""
    with _lock:
        f = _fifos.pop(name, None)
    if f is not None:
        if abort:
            f.abort()
        else:
            f.close()


def abort_fifo(name: str) -> None:
    with _lock:
        f = _fifos.get(name)
    if f is not None:
        f.abort()


def port_order(nports: int, keep_port_order: bool, seed: int) -> List[int]:
    """The order in which a vertex visits its input ports (see VertexReader)."""
    order = list(range(nports))
    if not keep_port_order and nports > 1:
        import random
        random.Random(seed).shuffle(order)
    return order
