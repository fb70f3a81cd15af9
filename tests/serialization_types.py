"""User record types for the serialization tests (module level, so other processes can import them) — the
shapes of the reference's SerializationTestTypes.cs: plain classes, nesting, arrays, enums, slots, frozen."""
import dataclasses
import enum


class Color(enum.Enum):
    RED = 1
    GREEN = "g"


@dataclasses.dataclass
class Point:
    x: int
    y: float


@dataclasses.dataclass(frozen=True)
class Frozen:
    name: str
    tags: tuple


@dataclasses.dataclass
class Nested:
    p: Point
    pts: list
    color: Color
    maybe: object = None


class Plain:
    def __init__(self, a=0, b=None):
        self.a = a
        self.b = b if b is not None else []

    def __eq__(self, other):
        return type(other) is Plain and (self.a, self.b) == (other.a, other.b)

    def __hash__(self):
        return hash(self.a)


class Slotted:
    __slots__ = ("u", "v")

    def __init__(self, u, v):
        self.u, self.v = u, v

    def __eq__(self, other):
        return type(other) is Slotted and (self.u, self.v) == (other.u, other.v)

    def __hash__(self):
        return hash((self.u, self.v))


class Derived(Plain):
    def __init__(self, a=0, b=None, c="c"):
        super().__init__(a, b)
        self.c = c

    def __eq__(self, other):
        return type(other) is Derived and (self.a, self.b, self.c) == (other.a, other.b, other.c)

    def __hash__(self):
        return hash((self.a, self.c))


class Node:
    def __init__(self, v):
        self.v = v
        self.next = None
