"""Small XML helpers of the plan writer — `DryadQueryDoc` (LinqToDryad/DryadLinqQueryDoc.cs:29-51):
element construction with text, attributes and CDATA sections (the per-vertex `<Explain>` text is
CDATA so that `<`, `&&` of rendered lambdas survive)."""
from __future__ import annotations

from typing import Iterable, Mapping, Optional
from xml.sax.saxutils import escape, quoteattr


class QueryDoc:
    def __init__(self, root: str = "Query"):
        self.root = root
        self._lines = [f"<{root}>"]
        self._depth = 1

    def _pad(self) -> str:
        return "  " * self._depth

    def element(self, name: str, text: Optional[object] = None, attrs: Optional[Mapping[str, object]] = None) -> "QueryDoc":
        a = "".join(f" {k}={quoteattr(str(v))}" for k, v in (attrs or {}).items())
        if text is None:
            self._lines.append(f"{self._pad()}<{name}{a} />")
        else:
            self._lines.append(f"{self._pad()}<{name}{a}>{escape(str(text))}</{name}>")
        return self

    def cdata(self, name: str, text: str) -> "QueryDoc":
        safe = text.replace("]]>", "]]]]><![CDATA[>")
        self._lines.append(f"{self._pad()}<{name}><![CDATA[{safe}]]></{name}>")
        return self

    def open(self, name: str, attrs: Optional[Mapping[str, object]] = None) -> "QueryDoc":
        a = "".join(f" {k}={quoteattr(str(v))}" for k, v in (attrs or {}).items())
        self._lines.append(f"{self._pad()}<{name}{a}>")
        self._depth += 1
        return self

    def close(self, name: str) -> "QueryDoc":
        self._depth -= 1
        self._lines.append(f"{self._pad()}</{name}>")
        return self

    def extend(self, lines: Iterable[str]) -> "QueryDoc":
        self._lines.extend(lines)
        return self

    def to_string(self) -> str:
        return "\n".join(self._lines + [f"</{self.root}>"]) + "\n"
