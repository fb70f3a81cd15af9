"""Fixed-width byte records (TeraSort-style 100-byte rows) in the query API.

`ctx.from_rows(tensor)` makes a dataset whose partitions are `[n, width]` uint8 tensors (pinned
host memory or HBM); `RowKey(offset, length)` is the key selector for such rows: the big-endian
byte string `row[offset:offset+length]`, compared ordinally — the same order as the reference's
LineRecord / byte[] keys (LinqToDryad/LineRecord.cs:144-154).
"""
from __future__ import annotations


class RowKey:
    """Key selector over byte rows; callable on host records (bytes) and recognised by the GPU
    OrderBy / RangePartition operators."""

    def __init__(self, offset: int = 0, length: int = 10):
        if length < 1:
            raise ValueError("RowKey length must be positive")
        self.offset, self.length = int(offset), int(length)
        self.__name__ = f"RowKey[{offset}:{offset + length}]"

    def __call__(self, row):
        return bytes(row[self.offset:self.offset + self.length])

    def __eq__(self, other):
        return isinstance(other, RowKey) and (self.offset, self.length) == (other.offset, other.length)

    def __hash__(self):
        return hash(("RowKey", self.offset, self.length))
