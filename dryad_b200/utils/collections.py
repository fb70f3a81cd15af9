"""Small utilities of the client library (reference: LinqToDryad/DryadLinqStringDictionary.cs,
DryadLinqStringList.cs — collections that become read-only once a job was submitted —
LinqToDryad/BitVector.cs)."""
from __future__ import annotations

from typing import Dict, Iterable, Iterator, List


class _Freezable:
    _frozen = False

    def freeze(self) -> None:
        self._frozen = True

    def _check(self):
        if self._frozen:
            from ..api.errors import DryadLinqException, ErrorCode
            raise DryadLinqException(ErrorCode.JobOptionNotImplemented, "collection is read-only after job submission")


class ReadOnlyAfterSubmitDict(_Freezable, dict):
    def __setitem__(self, k, v):
        self._check()
        super().__setitem__(k, v)

    def __delitem__(self, k):
        self._check()
        super().__delitem__(k)

    def update(self, *a, **k):
        self._check()
        super().update(*a, **k)

    def pop(self, *a):
        self._check()
        return super().pop(*a)

    def clear(self):
        self._check()
        super().clear()


class ReadOnlyAfterSubmitList(_Freezable, list):
    def append(self, v):
        self._check()
        super().append(v)

    def extend(self, v):
        self._check()
        super().extend(v)

    def remove(self, v):
        self._check()
        super().remove(v)

    def clear(self):
        self._check()
        super().clear()


class BitVector:
    """Growable bit set."""

    def __init__(self, nbits: int = 0):
        self._words: List[int] = [0] * ((nbits + 63) // 64)
        self._n = nbits

    def __len__(self):
        return self._n

    def set(self, i: int, value: bool = True) -> None:
        if i >= self._n:
            self._words.extend([0] * ((i + 64) // 64 - len(self._words)))
            self._n = i + 1
        if value:
            self._words[i >> 6] |= 1 << (i & 63)
        else:
            self._words[i >> 6] &= ~(1 << (i & 63))

    def __getitem__(self, i: int) -> bool:
        return i < self._n and bool((self._words[i >> 6] >> (i & 63)) & 1)

    def count(self) -> int:
        return sum(bin(w).count("1") for w in self._words)

    def __iter__(self) -> Iterator[int]:
        return (i for i in range(self._n) if self[i])
