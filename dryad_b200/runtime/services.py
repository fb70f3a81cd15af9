"""Native host services: message pump, versioned mailbox, delay scheduler (csrc/host/runtime.cpp);
references: GraphManager/kernel/DrMessagePump.h, ProcessService/ProcessService.cs:674-746,
LocalScheduler/LocalScheduler.cs."""
from __future__ import annotations

import ctypes as C
from typing import Callable, List, Optional, Sequence, Tuple

from .. import _native as N
from . import hostlib


class MessagePump:
    def __init__(self, handler: Callable[[int, int], None], threads: int = 4):
        self.lib = N.host()
        self._cb = hostlib.PUMP_HANDLER(lambda listener, message, _u: handler(listener, message))
        self.h = self.lib.pump_create(threads, self._cb, None)

    def post(self, listener: int, message: int, delay: float = 0.0) -> None:
        if self.h:
            self.lib.pump_post(self.h, listener, message, delay)

    def post_on_stream_completion(self, listener: int, message: int, stream: Optional[int] = None) -> None:
        """Post (listener, message) when all GPU work enqueued so far on `stream` (default: the
        current torch stream) has finished — cudaLaunchHostFunc, no thread blocks in a synchronize.
        This is how a vertex's kernels report their own completion to the event loop."""
        import torch
        st = torch.cuda.current_stream().cuda_stream if stream is None else stream
        arg = self.lib.pump_make_notifier(self.h, listener, message)
        N.check(N.kernels().dryad_stream_host_func(st, self.lib.pump_notifier_fn(), arg), "stream_host_func")

    def wait_delivered(self, count: int, timeout: float = 10.0) -> int:
        return self.lib.pump_wait_delivered(self.h, count, timeout)

    def close(self):
        if self.h:
            self.lib.pump_destroy(self.h)
            self.h = None


class Mailbox:
    """Versioned key/value store with long-poll semantics (the GM <-> worker control plane)."""

    def __init__(self):
        self.lib = N.host()
        self.h = self.lib.mbox_create()

    def set(self, key: str, status: str, data: bytes = b"") -> int:
        return self.lib.mbox_set(self.h, key.encode(), status.encode(), data, len(data))

    def block_on_status(self, key: str, last_seen: int = 0, timeout: float = 1.0) -> Optional[Tuple[int, str, bytes]]:
        st = C.create_string_buffer(256)
        cap = 1 << 16
        buf = C.create_string_buffer(cap)
        n = C.c_int64()
        v = self.lib.mbox_block_on_status(self.h, key.encode(), last_seen, timeout, st, 256, buf, cap, C.byref(n))
        if v == 0:
            return None
        while n.value > cap:                           # a value larger than the buffer: read it again, whole
            cap = int(n.value) + 1024
            buf = C.create_string_buffer(cap)
            v = self.lib.mbox_block_on_status(self.h, key.encode(), max(0, last_seen), 0.0, st, 256, buf, cap, C.byref(n))
            if v == 0:
                return None
        return v, st.value.decode(), buf.raw[:n.value]

    def close(self):
        if self.h:
            self.lib.mbox_destroy(self.h)
            self.h = None


class LocalScheduler:
    """Delay scheduling: computer -> rack (1 s) -> cluster (2 s); hard constraints never relax."""

    def __init__(self, rack_of: Sequence[int], rack_delay: float = 1.0, cluster_delay: float = 2.0):
        self.lib = N.host()
        n = len(rack_of)
        self.h = self.lib.ls_create(n, (C.c_int32 * n)(*rack_of), rack_delay, cluster_delay)

    def submit(self, proc_id: int, computers: Sequence[int] = (), hard: bool = False) -> None:
        n = len(computers)
        self.lib.ls_submit(self.h, proc_id, (C.c_int32 * max(n, 1))(*computers), n, int(hard))

    def computer_idle(self, computer: int, age_bias: float = 0.0) -> Optional[int]:
        r = self.lib.ls_computer_idle(self.h, computer, age_bias)
        return None if r < 0 else r

    @property
    def waiting(self) -> int:
        return self.lib.ls_waiting(self.h)

    def close(self):
        if self.h:
            self.lib.ls_destroy(self.h)
            self.h = None
