"""Block channel over device memory: the tensor face of the `dryad_chan_*` C ABI
(csrc/kernels/blockchan.cu; reference: wrappernativeinfo.h:103-150, LinqToDryad/NativeBlockStream.cs:34-235
`ReadDataBlock / WriteDataBlock / AllocateDataBlock / ReleaseDataBlock / Flush / Close / GetTotalLength`).

A `DeviceBlockChannel` streams byte blocks between a producer and a consumer that run on different
CUDA streams (typically different host threads): the pool is one device buffer cut into `nblocks`
blocks, so the producer is throttled to `nblocks` blocks ahead; blocks are handed over with CUDA
events, never with a host synchronisation of the producer.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterator, Optional

import torch

from .. import _native as N
from ..ops import core as K
from ..parallel.symm import tensor_from_ptr


class DeviceBlockChannel:
    def __init__(self, block_bytes: int = 1 << 20, nblocks: int = 4, device=None, checksum: bool = False,
                 buffer: Optional[torch.Tensor] = None, uri: str = ""):
        self.block_bytes = (int(block_bytes) + 15) // 16 * 16
        self.nblocks = int(nblocks)
        self.device = torch.device(device if device is not None else "cuda")
        if buffer is None:
            buffer = torch.empty(self.block_bytes * self.nblocks, dtype=torch.uint8, device=self.device)
        if buffer.numel() < self.block_bytes * self.nblocks:
            raise ValueError("buffer smaller than nblocks * block_bytes")
        self.buffer = buffer                                   # e.g. a slice of the symmetric arena
        lib = N.kernels()
        self.control = torch.zeros(lib.dryad_chan_control_bytes(self.nblocks), dtype=torch.uint8, device=buffer.device) if checksum else None
        h = C.c_void_p()
        N.check(lib.dryad_chan_create(K.ptr(buffer), self.block_bytes, self.nblocks, int(checksum), K.ptr(self.control), C.byref(h)),
                "chan_create")
        self.h = h
        self.uri = uri or f"hbm://{buffer.data_ptr():x}"
        self.timeout = 30.0

    # ---- writer ----------------------------------------------------------------------------------
    def alloc(self, nbytes: int) -> torch.Tensor:
        """AllocateDataBlock: a uint8 view of a free block (waits while the reader is nblocks behind)."""
        p = C.c_void_p()
        N.check(N.kernels().dryad_chan_alloc_block(self.h, nbytes, self.timeout, C.byref(p)), "chan_alloc_block")
        return tensor_from_ptr(p.value, self.block_bytes, self.buffer.device, owner=self.buffer)[:nbytes]

    def publish(self, block: torch.Tensor, nbytes: Optional[int] = None) -> None:
        """WriteDataBlock: hand a block obtained from alloc() to the reader (contents produced on the current stream)."""
        nb = block.numel() if nbytes is None else nbytes
        N.check(N.kernels().dryad_chan_write_block(self.h, block.data_ptr(), nb, K.stream()), "chan_write_block")

    def write(self, data: torch.Tensor) -> None:
        """Copy `data` (any dtype, contiguous) into the channel, split into blocks."""
        flat = data.contiguous().view(torch.uint8).reshape(-1)
        for a in range(0, max(flat.numel(), 1), self.block_bytes):
            chunk = flat[a:a + self.block_bytes]
            blk = self.alloc(chunk.numel())
            blk.copy_(chunk, non_blocking=True)
            self.publish(blk)
            if flat.numel() == 0:
                break

    def close(self) -> None:
        N.check(N.kernels().dryad_chan_close_writer(self.h), "chan_close_writer")

    # ---- reader ----------------------------------------------------------------------------------
    def read(self) -> Optional[torch.Tensor]:
        """ReadDataBlock: the next block as a uint8 view valid on the CURRENT stream (None at end of
        stream).  Call release(block) when the work that reads it has been queued."""
        p, n = C.c_void_p(), C.c_size_t()
        rc = N.kernels().dryad_chan_read_block(self.h, C.byref(p), C.byref(n), self.timeout, K.stream())
        if rc == 1:
            return None
        N.check(rc, "chan_read_block")
        return tensor_from_ptr(p.value, self.block_bytes, self.buffer.device, owner=self.buffer)[:n.value]

    def release(self, block: torch.Tensor) -> None:
        N.check(N.kernels().dryad_chan_release_block(self.h, block.data_ptr(), K.stream()), "chan_release_block")

    def __iter__(self) -> Iterator[torch.Tensor]:
        while True:
            b = self.read()
            if b is None:
                return
            yield b

    # ---- bookkeeping -----------------------------------------------------------------------------
    @property
    def total_length(self) -> int:
        return int(N.kernels().dryad_chan_total_length(self.h))

    def stats(self) -> dict:
        out = (C.c_int64 * 5)()
        N.check(N.kernels().dryad_chan_stats(self.h, out), "chan_stats")
        return {"blocks_written": out[0], "blocks_read": out[1], "bytes": out[2], "throttle_waits": out[3],
                "checksum_errors": out[4]}

    def destroy(self) -> None:
        if self.h:
            N.kernels().dryad_chan_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:  # noqa: BLE001
            pass


_registry: dict = {}


def open_device_channel(uri: str, **kw) -> DeviceBlockChannel:
    """`hbm://<name>[?blocks=N&block=BYTES&checksum=1]`: writer and reader of one process (two
    vertices of a start clique on two streams / threads) open the SAME channel by name."""
    import urllib.parse
    u = urllib.parse.urlparse(uri)
    if u.scheme != "hbm":
        raise ValueError(f"not a device channel uri: {uri}")
    name = u.netloc + u.path
    ch = _registry.get(name)
    if ch is None:
        q = urllib.parse.parse_qs(u.query)
        ch = DeviceBlockChannel(block_bytes=int(q.get("block", [kw.get("block_bytes", 1 << 20)])[0]),
                                nblocks=int(q.get("blocks", [kw.get("nblocks", 4)])[0]),
                                checksum=bool(int(q.get("checksum", ["0"])[0])), device=kw.get("device"), uri=uri)
        _registry[name] = ch
    return ch


def close_device_channel(uri: str) -> None:
    import urllib.parse
    u = urllib.parse.urlparse(uri)
    ch = _registry.pop(u.netloc + u.path, None)
    if ch is not None:
        ch.destroy()
