"""Rank-to-rank streamed channel (C7): a pointwise FIFO / pipe edge between two stages that live on
different ranks.

Reference: pointwise `DCT_Fifo` / `DCT_Pipe` edges (DryadLinqGraphManager/GraphBuilder.cs:533-540) —
the producer's item arrays flow to the consumer while both run (32-deep FIFO,
GraphManager/vertex/DrOutputGenerator.cpp:237), which is why both belong to one start clique.

Here a block is a tensor (one column chunk of a ColumnBatch, or a uint8 block of serialised
records).  Blocks travel with `torch.distributed` point-to-point operations (NCCL between GPUs, gloo
between CPU ranks) issued on a SIDE stream, so the transfer of block i overlaps the producer's
compute of block i+1 and the consumer's compute on block i-1; at most `depth` sends are in flight
(back-pressure, like the bounded FIFO).  A header tensor (kind, dtype code, numel) precedes every
block; numel < 0 closes the stream.
"""
from __future__ import annotations

from collections import deque
from typing import Deque, Iterator, Optional

import torch
import torch.distributed as dist

_DTYPES = [torch.uint8, torch.int8, torch.int16, torch.int32, torch.int64, torch.float16, torch.bfloat16,
           torch.float32, torch.float64, torch.bool]
_CODE = {dt: i for i, dt in enumerate(_DTYPES)}
PIPE_DEPTH = 32


class PipeWriter:
    def __init__(self, dst: int, device, tag: int = 0, depth: int = PIPE_DEPTH, group=None):
        self.dst, self.device, self.tag, self.depth, self.group = dst, torch.device(device), tag, depth, group
        self._cuda = self.device.type == "cuda"
        self._stream = torch.cuda.Stream(device=self.device) if self._cuda else None
        self._inflight: Deque = deque()
        self.blocks = 0
        self.bytes = 0

    def _send(self, t: torch.Tensor):
        return dist.isend(t, self.dst, group=self.group, tag=self.tag)

    def write(self, block: torch.Tensor) -> None:
        """Queue one block.  The tensor must stay untouched until it has been sent; the writer keeps
        a reference and waits for the oldest send when `depth` are outstanding."""
        block = block.contiguous()
        hdr = torch.tensor([1, _CODE[block.dtype], block.numel()], dtype=torch.int64, device=self.device)
        if self._cuda:
            self._stream.wait_stream(torch.cuda.current_stream(self.device))   # the block has been produced
            with torch.cuda.stream(self._stream):
                works = (self._send(hdr), self._send(block))
        else:
            works = (self._send(hdr), self._send(block))
        self._inflight.append((works, hdr, block))
        self.blocks += 1
        self.bytes += block.numel() * block.element_size()
        while len(self._inflight) > self.depth:
            self._retire()

    def _retire(self) -> None:
        works, _hdr, _blk = self._inflight.popleft()
        for w in works:
            w.wait()

    def close(self) -> None:
        hdr = torch.tensor([0, 0, -1], dtype=torch.int64, device=self.device)
        if self._cuda:
            self._stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self._stream):
                w = self._send(hdr)
        else:
            w = self._send(hdr)
        self._inflight.append(((w,), hdr, None))
        while self._inflight:
            self._retire()
        if self._cuda:
            torch.cuda.current_stream(self.device).wait_stream(self._stream)


class PipeReader:
    def __init__(self, src: int, device, tag: int = 0, group=None):
        self.src, self.device, self.tag, self.group = src, torch.device(device), tag, group
        self._cuda = self.device.type == "cuda"
        self._stream = torch.cuda.Stream(device=self.device) if self._cuda else None
        self.blocks = 0
        self.bytes = 0

    def _recv(self, t: torch.Tensor) -> None:
        dist.recv(t, self.src, group=self.group, tag=self.tag)

    def read(self) -> Optional[torch.Tensor]:
        """Next block, or None at end of stream.  The returned tensor is ready on the CURRENT stream."""
        hdr = torch.empty(3, dtype=torch.int64, device=self.device)
        if self._cuda:
            with torch.cuda.stream(self._stream):
                self._recv(hdr)
            self._stream.synchronize()                       # the header sizes the next receive
        else:
            self._recv(hdr)
        kind, code, numel = (int(x) for x in hdr.tolist())
        if numel < 0:
            return None
        blk = torch.empty(numel, dtype=_DTYPES[code], device=self.device)
        if self._cuda:
            with torch.cuda.stream(self._stream):
                self._recv(blk)
            torch.cuda.current_stream(self.device).wait_stream(self._stream)
        else:
            self._recv(blk)
        self.blocks += 1
        self.bytes += numel * blk.element_size()
        return blk

    def __iter__(self) -> Iterator[torch.Tensor]:
        while True:
            b = self.read()
            if b is None:
                return
            yield b


def send_batch(writer: PipeWriter, batch, chunk_rows: int = 1 << 20) -> None:
    """Stream a ColumnBatch column by column in row chunks (names and dtypes are known to both
    ends from the plan's schema, so only the data travels)."""
    names = sorted(batch.cols)
    n = batch.n
    for a in range(0, max(n, 1), chunk_rows):
        b = min(n, a + chunk_rows)
        for k in names:
            writer.write(batch.cols[k][a:b].reshape(-1))
        if n == 0:
            break
    writer.close()


def recv_batch(reader: PipeReader, names, inner=None):
    """Inverse of send_batch: returns dict name -> tensor (chunks concatenated)."""
    names = sorted(names)
    parts = {k: [] for k in names}
    i = 0
    for blk in reader:
        parts[names[i % len(names)]].append(blk)
        i += 1
    out = {}
    for k in names:
        t = torch.cat(parts[k]) if parts[k] else torch.empty(0, device=reader.device)
        if inner and inner.get(k, 1) > 1:
            t = t.view(-1, inner[k])
        out[k] = t
    return out
