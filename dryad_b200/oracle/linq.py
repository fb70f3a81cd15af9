"""LINQ-to-Objects semantics of every operator: the LocalDebug execution mode and the oracle for
differential tests (reference: LinqToDryad/DryadLinqEnumerable.cs:39-1423 and the
System.Linq.Enumerable behaviour DryadLinqLocalProvider delegates to,
LinqToDryad/DryadLinqQuery.cs:54-134).

All functions take and return plain Python lists (one logical sequence — in LocalDebug the
reference treats HashPartition / RangePartition as identity, DryadLinqEnumerable.cs:41-168).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence, Tuple

INT64_MIN, INT64_MAX = -(1 << 63), (1 << 63) - 1


class Grouping(list):
    """IGrouping<K, T>: a list of elements with a `.key`."""

    def __init__(self, key, items=()):
        super().__init__(items)
        self.key = key

    def __repr__(self):
        return f"Grouping({self.key!r}, {list(self)!r})"

    def __eq__(self, other):
        return isinstance(other, Grouping) and self.key == other.key and list(self) == list(other)

    def __hash__(self):
        return hash((self.key, tuple(self)))

    # LINQ-style aggregates over the group (so result selectors work on materialised groups too)
    Key = property(lambda s: s.key)

    def count(self, pred=None):
        return sum(1 for x in self if pred is None or pred(x))

    def sum(self, sel=None):
        return sum_(list(self), sel)

    def min(self, sel=None):
        return min_(list(self), sel)

    def max(self, sel=None):
        return max_(list(self), sel)

    def average(self, sel=None):
        return average(list(self), sel)

    def any(self, pred=None):
        return any_(list(self), pred)

    def all(self, pred):
        return all_(list(self), pred)

    def first(self):
        return first(list(self))

    def last(self):
        return last(list(self))

    def aggregate(self, decomposer, sel=None):
        inst = decomposer() if isinstance(decomposer, type) else decomposer
        inst.initialize(None)
        acc = None
        for i, x in enumerate(self):
            v = sel(x) if sel else x
            acc = inst.seed(v) if i == 0 else inst.accumulate(acc, v)
        return inst.final_reduce(acc)

    Count, Sum, Min, Max, Average, Any, All, First, Last, Aggregate = count, sum, min, max, average, any, all, first, last, aggregate


def _hashable(k):
    if isinstance(k, dict):
        return tuple(sorted((kk, _hashable(v)) for kk, v in k.items()))
    if isinstance(k, list):
        return tuple(_hashable(v) for v in k)
    return k


def checked_add(a, b):
    """C# `checked` arithmetic for integral sums (DryadLinqDecomposition.cs:312-315)."""
    r = a + b
    if isinstance(r, int) and not isinstance(r, bool) and not (INT64_MIN <= r <= INT64_MAX):
        raise OverflowError("Arithmetic operation resulted in an overflow.")
    return r


# ---- element-wise ------------------------------------------------------------------------------
def select(src: Sequence, fn: Callable, indexed: bool = False) -> List:
    return [fn(x, i) for i, x in enumerate(src)] if indexed else [fn(x) for x in src]


def where(src: Sequence, pred: Callable, indexed: bool = False) -> List:
    return [x for i, x in enumerate(src) if pred(x, i)] if indexed else [x for x in src if pred(x)]


def select_many(src: Sequence, fn: Callable, result: Optional[Callable] = None, indexed: bool = False) -> List:
    if not indexed and result is None:
        from itertools import chain
        return list(chain.from_iterable(map(fn, src)))     # C-speed flattening for the common form
    out = []
    for i, x in enumerate(src):
        for y in (fn(x, i) if indexed else fn(x)):
            out.append(result(x, y) if result else y)
    return out


def take(src, n): return list(src[:max(n, 0)])
def skip(src, n): return list(src[max(n, 0):])


def take_while(src, pred, indexed=False):
    out = []
    for i, x in enumerate(src):
        if not (pred(x, i) if indexed else pred(x)):
            break
        out.append(x)
    return out


def skip_while(src, pred, indexed=False):
    i = 0
    for i, x in enumerate(src):
        if not (pred(x, i) if indexed else pred(x)):
            return list(src[i:])
    return []


def concat(a, b): return list(a) + list(b)
def reverse(src): return list(reversed(src))
def zip_(a, b, fn): return [fn(x, y) for x, y in zip(a, b)]


# ---- ordering ---------------------------------------------------------------------------------
def order_by(src: Sequence, key: Callable, descending: bool = False, comparer=None) -> List:
    """Stable sort (Enumerable.OrderBy is stable; descending keeps the original order of ties)."""
    if comparer is not None:
        import functools
        cmpf = (lambda a, b: -comparer(key(a), key(b))) if descending else (lambda a, b: comparer(key(a), key(b)))
        return sorted(src, key=functools.cmp_to_key(cmpf))
    if not descending:
        return sorted(src, key=key)
    dec = sorted(((key(x), i, x) for i, x in enumerate(src)), key=lambda t: _Rev(t[0], t[1]))
    return [t[2] for t in dec]


class _Rev:
    """Sort helper: descending on the key, ascending on the original index (stability)."""
    __slots__ = ("k", "i")

    def __init__(self, k, i):
        self.k, self.i = k, i

    def __lt__(self, o):
        if self.k == o.k:
            return self.i < o.i
        return self.k > o.k


# ---- grouping ---------------------------------------------------------------------------------
def group_by(src: Sequence, key: Callable, elem: Optional[Callable] = None,
             result: Optional[Callable] = None) -> List:
    groups: "OrderedDict[Any, Grouping]" = OrderedDict()
    for x in src:
        k = key(x)
        hk = _hashable(k)
        g = groups.get(hk)
        if g is None:
            g = groups[hk] = Grouping(k)
        g.append(elem(x) if elem else x)
    if result:
        return [result(g.key, g) for g in groups.values()]
    return list(groups.values())


def distinct(src: Sequence) -> List:
    seen, out = set(), []
    for x in src:
        h = _hashable(x)
        if h not in seen:
            seen.add(h)
            out.append(x)
    return out


def union(a, b): return distinct(list(a) + list(b))


def intersect(a, b):
    sb = {_hashable(x) for x in b}
    return [x for x in distinct(a) if _hashable(x) in sb]


def except_(a, b):
    sb = {_hashable(x) for x in b}
    return [x for x in distinct(a) if _hashable(x) not in sb]


# ---- joins ------------------------------------------------------------------------------------
def join(outer, inner, okey, ikey, result):
    lookup: Dict[Any, List] = {}
    for y in inner:
        lookup.setdefault(_hashable(ikey(y)), []).append(y)
    out = []
    for x in outer:
        for y in lookup.get(_hashable(okey(x)), ()):
            out.append(result(x, y))
    return out


def group_join(outer, inner, okey, ikey, result):
    lookup: Dict[Any, List] = {}
    for y in inner:
        lookup.setdefault(_hashable(ikey(y)), []).append(y)
    # the group is an IEnumerable with the LINQ aggregates (g.count(), g.sum(sel), ...), like a GroupBy group
    return [result(x, Grouping(okey(x), lookup.get(_hashable(okey(x)), ()))) for x in outer]


# ---- aggregates -------------------------------------------------------------------------------
def count(src, pred=None): return sum(1 for x in src if pred is None or pred(x))
def any_(src, pred=None): return any(True for x in src if pred is None or pred(x))
def all_(src, pred): return all(pred(x) for x in src)
def contains(src, v): return any(x == v for x in src)


def sum_(src, sel=None):
    acc = 0
    for x in src:
        acc = checked_add(acc, sel(x) if sel else x)
    return acc


def _nonempty(src, what):
    if len(src) == 0:
        raise ValueError(f"Sequence contains no elements ({what})")


def min_(src, sel=None):
    vals = [sel(x) if sel else x for x in src]
    _nonempty(vals, "Min")
    return min(vals)


def max_(src, sel=None):
    vals = [sel(x) if sel else x for x in src]
    _nonempty(vals, "Max")
    return max(vals)


def average(src, sel=None):
    """Average carries (sum, count): ints widen to long, result is double
    (DryadLinqDecomposition.cs:494-578)."""
    vals = [sel(x) if sel else x for x in src]
    _nonempty(vals, "Average")
    s = 0
    for v in vals:
        s = checked_add(s, v) if isinstance(v, int) else s + v
    return s / len(vals)


def aggregate(src, func, seed=None, has_seed=False, result=None):
    it = iter(src)
    if has_seed:
        acc = seed
    else:
        try:
            acc = next(it)
        except StopIteration:
            raise ValueError("Sequence contains no elements (Aggregate)")
    for x in it:
        acc = func(acc, x)
    return result(acc) if result else acc


def first(src, pred=None, default=False):
    for x in src:
        if pred is None or pred(x):
            return x
    if default:
        return None
    raise ValueError("Sequence contains no matching element (First)")


def last(src, pred=None, default=False):
    for x in reversed(src):
        if pred is None or pred(x):
            return x
    if default:
        return None
    raise ValueError("Sequence contains no matching element (Last)")


def single(src, pred=None, default=False):
    m = [x for x in src if pred is None or pred(x)]
    if len(m) == 1:
        return m[0]
    if len(m) == 0 and default:
        return None
    raise ValueError("Sequence contains %s matching element (Single)" % ("no" if not m else "more than one"))


def sequence_equal(a, b): return list(a) == list(b)


# ---- DryadLINQ-specific -----------------------------------------------------------------------
def sliding_window(src: Sequence, fn: Callable, w: int) -> List:
    """fn is applied to every window of w consecutive records (DryadLinqEnumerable.cs:188)."""
    return [fn(list(src[i:i + w])) for i in range(0, len(src) - w + 1)] if w > 0 else []


def apply(src: Sequence, fn: Callable) -> List:
    return list(fn(list(src)))


def apply2(a: Sequence, b: Sequence, fn: Callable) -> List:
    return list(fn(list(a), list(b)))


def fork(src: Sequence, fn: Callable, n: int) -> Tuple[List, ...]:
    """Mapper returns a tuple of optional values (None = absent), one per branch
    (ForkTuple semantics, LinqToDryad/ForkTuple.cs:43-367)."""
    outs: Tuple[List, ...] = tuple([] for _ in range(n))
    for x in src:
        t = fn(x)
        for i in range(n):
            if t[i] is not None:
                outs[i].append(t[i])
    return outs


def offsets(counts: Sequence[int]) -> List[int]:
    """Exclusive prefix sum of per-partition counts (DryadLinqEnumerable.Offsets :1131)."""
    out, acc = [], 0
    for c in counts:
        out.append(acc)
        acc += c
    return out
