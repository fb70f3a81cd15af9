"""Static dataset properties propagated through the plan: partitioning, ordering, distinctness.

Reference: LinqToDryad/DataSetInfo.cs:38-799 (PartitionInfo Random/Hash/Range, OrderByInfo,
DistinctInfo, DataSetInfo).  These are what let the planner SKIP shuffles and sorts when the data is
already suitably partitioned / ordered (`IsPartitionedBy`, `IsSamePartition`, `Rewrite`).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, List, Optional, Sequence

from ..expr.ir import Expr


class KeySelector:
    """A key-selector lambda plus its traced Expr (when traceable).  Equality of two selectors is
    structural on the Expr (ExpressionMatcher role) and falls back to function identity."""

    def __init__(self, fn: Optional[Callable], ir: Optional[Expr] = None, name: str = ""):
        self.fn, self.ir, self.name = fn, ir, name

    def matches(self, other: Optional["KeySelector"]) -> bool:
        if other is None:
            return False
        if self.ir is not None and other.ir is not None:
            return self.ir.same_as(other.ir)
        return self.fn is not None and self.fn is other.fn

    def rewrite(self, projection: Optional[dict]) -> Optional["KeySelector"]:
        """Push this selector through a projection `new_col -> Expr(old cols)`: succeeds when some
        output column IS this key expression (DataSetInfo.Rewrite, DataSetInfo.cs:296-332)."""
        if self.ir is None or projection is None:
            return None
        from ..expr.ir import col
        for out_name, e in projection.items():
            if isinstance(e, Expr) and e.same_as(self.ir):
                return KeySelector(None, col(out_name, e.kind), out_name)
        return None

    def __repr__(self):
        return f"Key({self.ir if self.ir is not None else getattr(self.fn, '__name__', 'λ')})"


IDENTITY_KEY = KeySelector(None, None, "identity")


class PartitionInfo:
    kind = "random"

    def __init__(self, count: int):
        self.count = count

    def is_partitioned_by(self, key: KeySelector, comparer=None) -> bool:
        return False

    def is_same_partition(self, other: "PartitionInfo") -> bool:
        return False

    def rewrite(self, projection) -> "PartitionInfo":
        return self

    def with_count(self, n: int) -> "PartitionInfo":
        return RandomPartition(n)

    def describe(self) -> str:
        return f"Random[{self.count}]"


class RandomPartition(PartitionInfo):
    kind = "random"


class HashPartition(PartitionInfo):
    kind = "hash"

    def __init__(self, key: KeySelector, count: int, comparer=None):
        super().__init__(count)
        self.key, self.comparer = key, comparer

    def is_partitioned_by(self, key, comparer=None):
        return self.key.matches(key) and self.comparer is comparer

    def is_same_partition(self, other):
        return (isinstance(other, HashPartition) and self.count == other.count
                and self.comparer is other.comparer)

    def rewrite(self, projection):
        k = self.key.rewrite(projection)
        return HashPartition(k, self.count, self.comparer) if k is not None else RandomPartition(self.count)

    def with_count(self, n):
        return RandomPartition(n)

    def describe(self):
        return f"Hash[{self.count}]({self.key})"


class RangePartition(PartitionInfo):
    kind = "range"

    def __init__(self, key: KeySelector, count: int, separators: Optional[Sequence] = None,
                 descending: bool = False, comparer=None):
        super().__init__(count)
        self.key, self.separators, self.descending, self.comparer = key, separators, descending, comparer

    def is_partitioned_by(self, key, comparer=None):
        return self.key.matches(key) and self.comparer is comparer

    def has_keys(self) -> bool:
        return self.separators is not None

    def is_same_partition(self, other):
        return (isinstance(other, RangePartition) and self.count == other.count
                and self.descending == other.descending and self.comparer is other.comparer
                and self.separators is not None and other.separators is not None
                and list(self.separators) == list(other.separators))

    def rewrite(self, projection):
        k = self.key.rewrite(projection)
        if k is None:
            return RandomPartition(self.count)
        return RangePartition(k, self.count, self.separators, self.descending, self.comparer)

    def describe(self):
        return f"Range[{self.count}{' desc' if self.descending else ''}]({self.key})"


@dataclass
class OrderByInfo:
    key: Optional[KeySelector] = None
    descending: bool = False
    comparer: Any = None

    @property
    def is_ordered(self) -> bool:
        return self.key is not None

    def is_ordered_by(self, key: KeySelector, comparer=None, descending: Optional[bool] = None) -> bool:
        if self.key is None or not self.key.matches(key) or self.comparer is not comparer:
            return False
        return descending is None or descending == self.descending

    def rewrite(self, projection) -> "OrderByInfo":
        if self.key is None:
            return self
        k = self.key.rewrite(projection)
        return OrderByInfo(k, self.descending, self.comparer) if k is not None else OrderByInfo()

    def describe(self) -> str:
        return f"OrderBy({self.key}{' desc' if self.descending else ''})" if self.key else "Unordered"


@dataclass
class DistinctInfo:
    is_distinct: bool = False
    comparer: Any = None


@dataclass
class DataSetInfo:
    partition: PartitionInfo = field(default_factory=lambda: RandomPartition(1))
    order: OrderByInfo = field(default_factory=OrderByInfo)
    distinct: DistinctInfo = field(default_factory=DistinctInfo)

    def rewrite(self, projection) -> "DataSetInfo":
        return DataSetInfo(self.partition.rewrite(projection), self.order.rewrite(projection), DistinctInfo())

    def keep_partition_only(self) -> "DataSetInfo":
        return DataSetInfo(self.partition, OrderByInfo(), DistinctInfo())

    def describe(self) -> str:
        return f"{self.partition.describe()}; {self.order.describe()}" + ("; Distinct" if self.distinct.is_distinct else "")


def compute_is_descending(keys: Sequence, comparer=None) -> Optional[bool]:
    """Infer direction from explicit range keys; inconsistent keys are an error
    (LinqToDryad/DataSetInfo.cs:250-276, DryadLinqUtil.ComputeIsDescending :523)."""
    import functools
    cmpf = comparer if comparer is not None else (lambda a, b: (a > b) - (a < b))
    direction = 0
    for a, b in zip(keys, keys[1:]):
        c = cmpf(a, b)
        if c == 0:
            continue
        d = -1 if c < 0 else 1
        if direction == 0:
            direction = d
        elif d != direction:
            from ..api.errors import DryadLinqException, ErrorCode
            raise DryadLinqException(ErrorCode.PartitionKeysAreNotConsistentlyOrdered,
                                     "range partition keys are not consistently ordered")
    return direction > 0 if direction != 0 else False
