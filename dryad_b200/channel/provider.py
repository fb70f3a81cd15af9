"""Storage back ends behind one interface, with a registration point for new ones.

Reference: LinqToDryad/DataProvider.cs:62-290 (`DataProvider`: Scheme, GetMetaData, GetStreamInfo,
GetTemporaryStreamUri, RewriteUri, Ingress, Egress, CheckExistence, static `Register` /
`GetDataProvider`), DryadLinqStreamInfo (:268-290), LinqToDryad/DataPath.cs:36-124.

`channel/store.py` routes every table operation through the provider registered for the URI's
scheme.  Built in: `partfile` (partitioned block-stream files) and `file` (plain text); a new scheme
is one subclass + `DataProvider.register(MyProvider())`.  `MemoryDataProvider` ("mem://name") is a
complete small provider: tables that live in the client process (useful in tests and notebooks).
"""
from __future__ import annotations

import abc
import threading
from dataclasses import dataclass
from typing import Any, Dict, List, Optional, Sequence


@dataclass
class StreamInfo:
    """Partition count and total stored size of a data set (DryadLinqStreamInfo)."""
    partition_count: int
    data_size: int


class DataProvider(abc.ABC):
    _providers: Dict[str, "DataProvider"] = {}
    _lock = threading.Lock()

    # ---- what a provider implements -----------------------------------------------------------
    @property
    @abc.abstractmethod
    def scheme(self) -> str: ...

    @abc.abstractmethod
    def exists(self, uri: str) -> bool: ...

    @abc.abstractmethod
    def delete(self, uri: str) -> None: ...

    @abc.abstractmethod
    def open_table(self, uri: str, text_partitions: int = 1):
        """-> object with `.partitions` (list of record lists / raw partitions) and `.schema`."""

    @abc.abstractmethod
    def write_table(self, uri: str, parts: Sequence[Sequence], schema=None, compression: Optional[str] = None) -> None:
        """Write all partitions and publish the table atomically."""

    def get_stream_info(self, uri: str) -> StreamInfo:
        t = self.open_table(uri)
        return StreamInfo(len(t.partitions), sum(len(p) for p in t.partitions))

    def get_metadata(self, uri: str) -> Dict[str, Any]:
        """Schema / compression record kept next to the data (DryadLinqMetaData)."""
        return {"schema": getattr(self.open_table(uri), "schema", None), "compression": "None"}

    def temporary_uri(self, context, path: str) -> str:
        """Where an unnamed (temporary) output of a job goes (GetTemporaryStreamUri)."""
        return f"{self.scheme}://{path}"

    def rewrite_uri(self, context, uri: str) -> str:
        return uri

    path_separator = "/"

    # distributed commit (one partition per writer, published once all are there).  Providers that
    # cannot do it leave these alone: the executors then gather the partitions and call write_table.
    supports_part_commit = False

    def write_part(self, uri: str, index: int, part, compression: Optional[str] = None):
        raise NotImplementedError

    def commit_part(self, tmp: str, final: str) -> None:
        raise NotImplementedError

    def finish_table(self, uri: str, sizes: Sequence[int], schema=None, compression: Optional[str] = None) -> None:
        raise NotImplementedError

    # ---- registry -------------------------------------------------------------------------------
    @staticmethod
    def register(provider: "DataProvider", replace: bool = False) -> None:
        from ..api.errors import DryadLinqException, ErrorCode
        with DataProvider._lock:
            if provider.scheme in DataProvider._providers and not replace:
                raise DryadLinqException(ErrorCode.UnrecognizedDataSource,
                                         f"a data provider for scheme {provider.scheme!r} is already registered")
            DataProvider._providers[provider.scheme] = provider

    @staticmethod
    def unregister(scheme: str) -> None:
        with DataProvider._lock:
            DataProvider._providers.pop(scheme, None)

    @staticmethod
    def get(scheme: str) -> "DataProvider":
        from ..api.errors import DryadLinqException, ErrorCode
        p = DataProvider._providers.get(scheme)
        if p is None:
            known = ", ".join(sorted(DataProvider._providers))
            hint = ("needs a cluster file system; register a DataProvider for it" if scheme in ("hdfs", "wasb", "azureblob", "hpcdsc")
                    else f"registered schemes: {known}")
            raise DryadLinqException(ErrorCode.UnrecognizedDataSource, f"unknown storage scheme {scheme!r} ({hint})")
        return p

    @staticmethod
    def for_uri(uri: str) -> "DataProvider":
        return DataProvider.get(uri.split("://", 1)[0] if "://" in uri else "partfile")

    @staticmethod
    def schemes() -> List[str]:
        return sorted(DataProvider._providers)


class _MemTable:
    def __init__(self, partitions, schema):
        self.partitions, self.schema = partitions, schema


class MemoryDataProvider(DataProvider):
    """`mem://name`: tables held in this process.  Publication is one dictionary assignment, so a
    failed job leaves nothing behind, like the file providers' rename commit."""

    def __init__(self):
        self.tables: Dict[str, _MemTable] = {}

    scheme = "mem"

    def exists(self, uri):
        return uri in self.tables

    def delete(self, uri):
        self.tables.pop(uri, None)

    def open_table(self, uri, text_partitions=1):
        from ..api.errors import DryadLinqException, ErrorCode
        t = self.tables.get(uri)
        if t is None:
            raise DryadLinqException(ErrorCode.StreamDoesNotExist, f"{uri} does not exist")
        return _MemTable([list(p) for p in t.partitions], t.schema)

    def write_table(self, uri, parts, schema=None, compression=None):
        staged = [list(p.to_records() if hasattr(p, "to_records") else p) for p in parts]
        self.tables[uri] = _MemTable(staged, schema)

    def get_stream_info(self, uri):
        t = self.open_table(uri)
        return StreamInfo(len(t.partitions), sum(len(p) for p in t.partitions))
