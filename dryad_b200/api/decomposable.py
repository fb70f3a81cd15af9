"""User-defined decomposable and associative aggregation protocols.

Reference: LinqToDryad/IDecomposable.cs:28-139 (Initialize, Seed, Accumulate, RecursiveAccumulate,
FinalReduce), LinqToDryad/IAssociative.cs:28-80 (Seed, RecursiveAccumulate), attributes
[Decomposable(type)] / [Associative(type)] (LinqToDryad/Attributes.cs:37-275).
"""
from __future__ import annotations

from typing import Any, Callable, Generic, Iterable, Optional, TypeVar

S = TypeVar("S")
A = TypeVar("A")
R = TypeVar("R")


class IDecomposable(Generic[S, A, R]):
    """f(group) = final_reduce( recursive_accumulate*( accumulate*( seed(x0), x... ) ) )."""

    def initialize(self, state: Any = None) -> None:
        pass

    def seed(self, x: S) -> A:
        raise NotImplementedError

    def accumulate(self, acc: A, x: S) -> A:
        raise NotImplementedError

    def recursive_accumulate(self, acc: A, other: A) -> A:
        raise NotImplementedError

    def final_reduce(self, acc: A) -> R:
        return acc  # type: ignore[return-value]


class IAssociative(Generic[A]):
    def seed(self, x: Any) -> A:
        return x

    def recursive_accumulate(self, acc: A, other: A) -> A:
        raise NotImplementedError


def decomposable(decomposer_type: type):
    """[Decomposable(typeof(T))]: marks a function over a sequence as decomposable through
    `decomposer_type` (which must implement IDecomposable)."""
    if not (isinstance(decomposer_type, type) and issubclass(decomposer_type, IDecomposable)):
        from .errors import DryadLinqException, ErrorCode
        raise DryadLinqException(ErrorCode.DecomposerTypeDoesNotImplementInterface,
                                 f"{decomposer_type!r} does not implement IDecomposable")

    def mark(fn: Callable) -> Callable:
        fn.__dryad_decomposable__ = decomposer_type
        return fn
    return mark


def associative(assoc_type: Optional[type] = None):
    """[Associative]: marks a binary function as associative (and, as in the reference, assumed
    commutative — DryadLinqQueryNode.cs:2879-2880), optionally with an IAssociative type."""
    def mark(fn: Callable) -> Callable:
        if assoc_type is not None and not issubclass(assoc_type, IAssociative):
            from .errors import DryadLinqException, ErrorCode
            raise DryadLinqException(ErrorCode.AssociativeTypeDoesNotImplementInterface,
                                     f"{assoc_type!r} does not implement IAssociative")
        fn.__dryad_associative__ = assoc_type or True
        return fn
    return mark


def resource(is_stateful: bool = True, is_expensive: bool = False):
    """[Resource(IsStateful, IsExpensive)]: scheduling/pipelining hints on user functions."""
    def mark(fn: Callable) -> Callable:
        fn.__dryad_resource__ = {"is_stateful": is_stateful, "is_expensive": is_expensive}
        return fn
    return mark


def nullable(*fields: str):
    """[Nullable]: the named fields of a record class (or every field when none are named) may be
    None; the serializer writes a presence byte for them (LinqToDryad/Attributes.cs:37-60).  The
    self-describing record encoding here always carries presence, so this only records intent —
    `nullable_fields(cls)` lets schema inference keep such columns off the GPU fast path."""
    def mark(cls):
        cls.__dryad_nullable__ = tuple(fields) or True
        return cls
    return mark


def nullable_fields(cls):
    return getattr(cls, "__dryad_nullable__", ())


def is_associative(fn: Callable) -> bool:
    import operator
    if getattr(fn, "__dryad_associative__", None):
        return True
    return fn in (operator.add, operator.mul, min, max, operator.and_, operator.or_, operator.xor)
