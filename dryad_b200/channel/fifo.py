"""Native channel primitives (csrc/host/channel.cpp): bounded FIFO of blocks, block cache with
spill, Rabin fingerprint.  References: channelfifo.cpp (.h:25-241), ProcessService/Cache.cs and
SpillMachine.cs, classlib/DrFPrint.cpp + LinqToDryad/Hash64.cs."""
from __future__ import annotations

import ctypes as C
from typing import Optional

from .. import _native as N

FIFO_DEPTH = 32          # fifo://32/... (GraphManager/vertex/DrOutputGenerator.cpp:237)


class Fifo:
    def __init__(self, depth: int = FIFO_DEPTH):
        self.lib = N.host()
        self.h = self.lib.fifo_create(depth)

    def write(self, block: bytes, timeout: float = 10.0) -> bool:
        rc = self.lib.fifo_write(self.h, block, len(block), timeout)
        if rc == -2:
            raise BrokenPipeError("fifo closed or aborted")
        return rc == 0

    def close_writer(self) -> None:
        self.lib.fifo_close_writer(self.h)

    def abort(self) -> None:
        self.lib.fifo_abort(self.h)

    def read(self, timeout: float = 10.0, cap: int = 1 << 22) -> Optional[bytes]:
        """A block, or None at end of stream.  Raises TimeoutError / BrokenPipeError."""
        buf = C.create_string_buffer(cap)
        n = self.lib.fifo_read(self.h, buf, cap, timeout)
        if n == -1:
            raise TimeoutError("fifo read timed out")
        if n == -2:
            return None
        if n == -3:
            raise BrokenPipeError("fifo aborted")
        if n == -4:
            raise ValueError("block larger than the read buffer")
        return buf.raw[:n]

    @property
    def depth_now(self) -> int:
        return self.lib.fifo_depth_now(self.h)

    def close(self):
        if self.h:
            self.lib.fifo_destroy(self.h)
            self.h = None


class BlockCache:
    def __init__(self, spill_dir: str, low_water: int, high_water: int):
        self.lib = N.host()
        self.h = self.lib.cache_create(spill_dir.encode(), low_water, high_water)

    def put(self, key: str, data: bytes) -> None:
        self.lib.cache_put(self.h, key.encode(), data, len(data))

    def get(self, key: str, cap: int = 1 << 24) -> Optional[bytes]:
        buf = C.create_string_buffer(cap)
        n = self.lib.cache_get(self.h, key.encode(), buf, cap)
        return None if n < 0 else buf.raw[:n]

    def stats(self) -> dict:
        a = [C.c_int64() for _ in range(4)]
        self.lib.cache_stats(self.h, *[C.byref(x) for x in a])
        return {"resident": a[0].value, "spills": a[1].value, "loads": a[2].value, "blocks": a[3].value}

    def close(self):
        if self.h:
            self.lib.cache_destroy(self.h)
            self.h = None


def fingerprint(data: bytes, seed: int = 0) -> int:
    return N.host().fprint64(seed, data, len(data))
