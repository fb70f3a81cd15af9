"""Text semantics: LineRecord line splitting (LinqToDryad/DryadLinqTextReader.cs) and the
WordCount sample's `Split(' ')` (samples/WordCount.cs.pp:72)."""
from __future__ import annotations

from typing import List


def split_lines(data: bytes) -> List[bytes]:
    if not data:
        return []
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    return [l[:-1] if l.endswith(b"\r") else l for l in lines]


def split_tokens(data: bytes, delim: bytes = b" ", keep_empty: bool = True) -> List[bytes]:
    """Tokens of every line.  keep_empty=True is String.Split(' ') (adjacent delimiters yield
    empty tokens); False keeps maximal non-delimiter runs only."""
    out: List[bytes] = []
    lines = data.split(b"\n")
    if lines and lines[-1] == b"":
        lines.pop()
    for line in lines:
        parts = line.split(delim)
        out.extend(parts if keep_empty else [p for p in parts if p])
    return out
