"""Vertex start commands and status reports on the native TLV wire format (csrc/host/command.cpp;
reference: dvertexcommand.h:48-243 DVertexCommandBlock / DVertexStatus, DrVertex::MakeVertexStartCommand,
GraphManager/vertex/DrVertex.cpp:356-429).  A launch descriptor travels GM -> worker through the
versioned mailbox; the worker answers with a status block carrying per-channel byte counts."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

from .. import _native as N

CMD_START, CMD_REOPEN_CHANNELS, CMD_TERMINATE = 1, 2, 3
STATE_RUNNING, STATE_COMPLETED, STATE_FAILED = 1, 2, 3


@dataclass
class VertexCommand:
    command: int = CMD_START
    vertex_id: int = 0
    version: int = 0
    args: List[str] = field(default_factory=list)
    inputs: List[Tuple[str, int]] = field(default_factory=list)      # (uri, expected length)
    outputs: List[Tuple[str, int]] = field(default_factory=list)     # (uri, size hint)
    state: int = 0
    exit_code: int = 0
    message: str = ""
    channel_status: List[Tuple[int, int]] = field(default_factory=list)   # (bytes processed, error code)

    def serialize(self) -> bytes:
        lib = N.host()
        h = lib.vc_create(self.command, self.vertex_id, self.version)
        try:
            for a in self.args:
                lib.vc_add_arg(h, a.encode())
            for u, l in self.inputs:
                lib.vc_add_input(h, u.encode(), l)
            for u, l in self.outputs:
                lib.vc_add_output(h, u.encode(), l)
            lib.vc_set_status(h, self.state, self.exit_code, self.message.encode())
            for b, e in self.channel_status:
                lib.vc_add_channel_status(h, b, e)
            n = lib.vc_serialize(h, None, 0)
            buf = C.create_string_buffer(n)
            lib.vc_serialize(h, buf, n)
            return buf.raw
        finally:
            lib.vc_destroy(h)

    @staticmethod
    def parse(data: bytes) -> Optional["VertexCommand"]:
        lib = N.host()
        h = lib.vc_parse(data, len(data))
        if not h:
            return None
        try:
            a = [C.c_int() for _ in range(5)]
            lib.vc_get_header(h, *[C.byref(x) for x in a])
            vc = VertexCommand(a[0].value, a[1].value, a[2].value, state=a[3].value, exit_code=a[4].value)
            cap = max(8192, len(data) + 16)            # no field is longer than the block it came in
            buf = C.create_string_buffer(cap)
            for i in range(lib.vc_count(h, 0)):
                lib.vc_get_item(h, 0, i, buf, cap)
                vc.args.append(buf.value.decode())
            for i in range(lib.vc_count(h, 1)):
                n = lib.vc_get_item(h, 1, i, buf, cap)
                vc.inputs.append((buf.value.decode(), n))
            for i in range(lib.vc_count(h, 2)):
                n = lib.vc_get_item(h, 2, i, buf, cap)
                vc.outputs.append((buf.value.decode(), n))
            for i in range(lib.vc_count(h, 3)):
                n = lib.vc_get_item(h, 3, i, buf, cap)
                vc.channel_status.append((n, int(buf.value.decode() or 0)))
            lib.vc_get_item(h, 4, 0, buf, cap)
            vc.message = buf.value.decode()
            return vc
        finally:
            lib.vc_destroy(h)


def terminate_command(commit: bool = True, message: str = "") -> VertexCommand:
    """The Terminate command of the vertex-host protocol (dvertexpncontrol.h:34-106)."""
    return VertexCommand(CMD_TERMINATE, 0, 0, args=["commit" if commit else "abort", message[:4000]])


def affinity_merge(weights: List[Tuple[int, int]], rack_of: List[int], threshold: float = 0.5) -> Tuple[str, int]:
    """(level, id): where a vertex should run given (computer, input bytes) affinities."""
    n = len(weights)
    comp = (C.c_int32 * max(n, 1))(*[c for c, _ in weights])
    w = (C.c_int64 * max(n, 1))(*[x for _, x in weights])
    racks = (C.c_int32 * max(len(rack_of), 1))(*rack_of)
    choice = C.c_int32()
    lvl = N.host().affinity_merge(n, comp, w, racks, len(rack_of), threshold, C.byref(choice))
    return ("computer", "rack", "cluster")[lvl], choice.value
