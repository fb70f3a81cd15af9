"""Record types of the public API (reference: LinqToDryad/LineRecord.cs:30-176, Pair in
LinqToDryad/DryadLinqExtension.cs:39-120, LinqToDryad/ForkTuple.cs:43-367, grouping types in
LinqToDryad/DryadLinqCollection.cs:137)."""
from __future__ import annotations

from typing import Any, Iterable, Optional

from ..oracle.linq import Grouping  # re-exported: IGrouping<K, T>


class Record(dict):
    """A named-field record: a dict with attribute access that is hashable and ordered by its
    field values (the analog of C# anonymous / user record types)."""

    __slots__ = ()

    def __getattr__(self, name):
        try:
            return self[name]
        except KeyError:
            raise AttributeError(name) from None

    def __setattr__(self, name, value):
        self[name] = value

    def __hash__(self):  # type: ignore[override]
        return hash(tuple(self.items()))

    def __lt__(self, other):
        return tuple(self.values()) < tuple(other.values())

    def __repr__(self):
        return "Record(" + ", ".join(f"{k}={v!r}" for k, v in self.items()) + ")"


class Pair(tuple):
    """Pair<K, V> with `.key` / `.value`."""

    def __new__(cls, key, value):
        return super().__new__(cls, (key, value))

    key = property(lambda s: s[0])
    value = property(lambda s: s[1])

    def __repr__(self):
        return f"Pair({self[0]!r}, {self[1]!r})"


class LineRecord(str):
    """A text line; compares ordinally (LineRecord.cs:144-154).  Python str compares by code point,
    which is ordinal comparison."""

    @property
    def line(self) -> str:
        return str(self)


class ForkTuple(tuple):
    """Result of a Fork mapper: element i is present when it is not None
    (HasFirst/HasSecond/HasThird, ForkTuple.cs:43-367)."""

    def __new__(cls, *items):
        return super().__new__(cls, items)

    first = property(lambda s: s[0])
    second = property(lambda s: s[1] if len(s) > 1 else None)
    third = property(lambda s: s[2] if len(s) > 2 else None)
    has_first = property(lambda s: s[0] is not None)
    has_second = property(lambda s: len(s) > 1 and s[1] is not None)
    has_third = property(lambda s: len(s) > 2 and s[2] is not None)


class AggregateValue:
    """(value, count) carrier used by *OrDefault aggregates (DryadLinqVertex.cs:5313)."""

    __slots__ = ("value", "count")

    def __init__(self, value: Any = None, count: int = 0):
        self.value, self.count = value, count


__all__ = ["Record", "Pair", "LineRecord", "ForkTuple", "AggregateValue", "Grouping"]


class ForkValue:
    """An element that may carry no value — the single-output sibling of ForkTuple
    (`ForkValue<T>`, LinqToDryad/ForkTuple.cs:32-110): `has_value` is false for a default
    instance, assigning `value` makes it valid; equal when both are empty or both values match;
    an empty instance hashes to 0."""

    __slots__ = ("_x", "_has")

    def __init__(self, *x):
        if len(x) > 1:
            raise TypeError("ForkValue takes at most one value")
        self._has = bool(x)
        self._x = x[0] if x else None

    @property
    def has_value(self) -> bool:
        return self._has

    @property
    def value(self):
        return self._x

    @value.setter
    def value(self, v):
        self._x, self._has = v, True

    HasValue = has_value
    Value = value

    def __eq__(self, o):
        if not isinstance(o, ForkValue):
            return NotImplemented
        return (not self._has and not o._has) or (self._has and o._has and self._x == o._x)

    def __hash__(self):
        return hash(self._x) if self._has else 0

    def __repr__(self):
        return f"ForkValue({self._x!r})" if self._has else "ForkValue()"
