"""Graph-manager option table and per-job application state — the role of `DryadLINQApp`
(DryadLinqGraphManager/DryadLinqApplication.cs:34-409): the knobs the dynamic managers read
(`-mai/-maxaggregateinputs` 150, `-mafi/-maxaggregatefilterinputs` 32, `-at/-aggregatethreshold`
1 GB with KB/MB/GB/TB/PB suffixes and fractions, `-nca/-noclusteraffinity`), the legacy no-op
binary-selection switches, and the identity-map file `<plan>.map` that records which plan vertex
became which graph-manager stage."""
from __future__ import annotations

import re
from dataclasses import dataclass, field
from pathlib import Path
from typing import Dict, List, Optional, Sequence, Tuple

_SUFFIX = {"": 1, "KB": 1 << 10, "MB": 1 << 20, "GB": 1 << 30, "TB": 1 << 40, "PB": 1 << 50}


def parse_data_size(text: str) -> int:
    """`12KB`, `20.5MB`, `1073741824`: bytes, case-insensitive suffix, fractions only with a suffix."""
    m = re.fullmatch(r"\s*([0-9]+(?:\.[0-9]+)?)\s*([KMGTP]B)?\s*", text, re.IGNORECASE)
    if not m:
        raise ValueError(f"bad data size {text!r}")
    num, suf = m.group(1), (m.group(2) or "").upper()
    if "." in num and not suf:
        raise ValueError(f"fractional data size needs a unit suffix: {text!r}")
    return int(float(num) * _SUFFIX[suf])


@dataclass(frozen=True)
class OptionDescription:
    category: str          # General / Inputs / Outputs / VertexBehavior
    short: str
    long: str
    arguments: str
    text: str


OPTIONS: Tuple[OptionDescription, ...] = (
    OptionDescription("General", "amd64", "useamd64binary", "", "Dummy argument for legacy reasons"),
    OptionDescription("General", "i386", "usei386binary", "", "Dummy argument for legacy reasons"),
    OptionDescription("General", "retail", "useretailbinary", "", "Dummy argument for legacy reasons"),
    OptionDescription("General", "debug", "usedebugbinary", "", "Dummy argument for legacy reasons"),
    OptionDescription("VertexBehavior", "mai", "maxaggregateinputs", "<maxInputs>",
                      "Only allow aggregate vertices to use up to maxInputs inputs."),
    OptionDescription("VertexBehavior", "mafi", "maxaggregatefilterinputs", "<maxInputs>",
                      "Only allow aggregate filter vertices to use up to maxInputs inputs."),
    OptionDescription("VertexBehavior", "at", "aggregatethreshold", "<dataSize>",
                      "Only allow aggregate and aggregate filter vertices to use inputs up to a total of "
                      "dataSize (bytes, or KB/MB/GB/TB/PB suffix; fractions allowed with a suffix)."),
    OptionDescription("VertexBehavior", "nca", "noclusteraffinity", "", "Ignore machine/pod affinity when grouping."),
)


@dataclass
class DryadLinqApp:
    max_aggregate_inputs: int = 150
    max_aggregate_filter_inputs: int = 32
    aggregate_threshold: int = 1 << 30
    cluster_affinity: bool = True
    positional: List[str] = field(default_factory=list)
    identity_map: Dict[int, int] = field(default_factory=dict)      # plan vertex id -> GM stage id
    _map_path: Optional[Path] = None

    @classmethod
    def parse(cls, argv: Sequence[str]) -> "DryadLinqApp":
        """Consume the GM's command-line options; unknown `-x` switches are an error, bare words
        (the plan file) are kept in `positional`."""
        app = cls()
        by_name = {}
        for o in OPTIONS:
            by_name[o.short] = by_name[o.long] = o
        i = 0
        argv = list(argv)
        while i < len(argv):
            a = argv[i]
            if not a.startswith("-"):
                app.positional.append(a)
                i += 1
                continue
            o = by_name.get(a.lstrip("-").lower())
            if o is None:
                raise ValueError(f"unknown graph manager option {a!r}")
            val = None
            if o.arguments:
                if i + 1 >= len(argv):
                    raise ValueError(f"option {a} needs {o.arguments}")
                val = argv[i + 1]
                i += 1
            if o.short == "mai":
                app.max_aggregate_inputs = int(val)
            elif o.short == "mafi":
                app.max_aggregate_filter_inputs = int(val)
            elif o.short == "at":
                app.aggregate_threshold = parse_data_size(val)
            elif o.short == "nca":
                app.cluster_affinity = False
            i += 1
        return app

    @staticmethod
    def usage() -> str:
        lines = []
        for o in OPTIONS:
            lines.append(f"  -{o.short:7s} -{o.long} {o.arguments}".rstrip())
            lines.append(f"           {o.text}")
        return "\n".join(lines)

    # ---- dynamic-manager parameters ---------------------------------------------------------
    def aggregate_groups(self, bytes_: List[int], machine: List[int], pod: List[int], filtered: bool = False) -> List[int]:
        """Runtime aggregation-tree grouping with this job's limits (filter vertices take fewer
        inputs; `-nca` collapses the topology so grouping ignores machine and pod)."""
        from .gm import aggregate_groups
        if not self.cluster_affinity:
            machine, pod = [0] * len(bytes_), [0] * len(bytes_)
        cap = self.max_aggregate_filter_inputs if filtered else self.max_aggregate_inputs
        return aggregate_groups(bytes_, machine, pod, max_inputs=cap, max_bytes=self.aggregate_threshold)

    # ---- identity map ------------------------------------------------------------------------
    def open_identity_map(self, plan_path: str) -> Path:
        self._map_path = Path(str(plan_path) + ".map")
        self._map_path.write_text("")
        return self._map_path

    def record_identity(self, plan_vertex: int, stage: int, name: str = "") -> None:
        self.identity_map[plan_vertex] = stage
        if self._map_path is not None:
            with self._map_path.open("a") as f:
                f.write(f"{plan_vertex}\t{stage}\t{name}\n")

    @staticmethod
    def read_identity_map(path) -> Dict[int, Tuple[int, str]]:
        out = {}
        for line in Path(path).read_text().splitlines():
            a, b, *rest = line.split("\t")
            out[int(a)] = (int(b), rest[0] if rest else "")
        return out
