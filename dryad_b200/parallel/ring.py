"""Device-side streamed channel between two GPUs: a ring of blocks in the consumer's HBM with
flag-based flow control (csrc/kernels/ring.cu).  The host only enqueues one push / pop kernel per
block; publication, back-pressure and slot recycling happen between the two GPUs.

`HbmRing(arena, src, dst, slots, slot_bytes)` is constructed on EVERY rank (symmetric allocation);
rank `src` may call `push`, rank `dst` may call `pop`.  Reference role: the bounded FIFO channel
between pipelined vertices (channelfifo.cpp) and the pipe/FIFO edge types of GraphBuilder.
"""
from __future__ import annotations

import torch

from .. import _native as N
from ..ops import core as K
from .symm import SymmetricArena


class HbmRing:
    def __init__(self, arena: SymmetricArena, src: int, dst: int, slots: int = 4, slot_bytes: int = 1 << 20):
        if src == dst:
            raise ValueError("HbmRing needs two different GPUs (kernels that wait on each other must not share one)")
        self.arena, self.src, self.dst = arena, src, dst
        self.slots, self.slot_bytes = int(slots), int((slot_bytes + 15) // 16 * 16)
        self.ring, self.ring_off = arena.alloc((self.slots * self.slot_bytes,), torch.uint8)
        self.ctl, self.ctl_off = arena.alloc((8 + self.slots,), torch.int64)   # [0] head  [1] tail  [2] done_push  [3] done_pop  [8..] lens
        self.ctl.zero_()
        arena.check_symmetric("HbmRing")
        arena.barrier()                                 # every rank's control words are zero before any push
        self.pushed = 0
        self.popped = 0
        self.dead = False                               # set by check(): a timed-out ring is never used again
        self.status = torch.zeros(1, dtype=torch.int32, device=arena.device)
        self.out_len = torch.zeros(1, dtype=torch.int64, device=arena.device)

    def _ctl(self, rank: int, word: int) -> int:
        return self.arena.peer_ptr(rank, self.ctl_off + 8 * word)

    def push(self, block: torch.Tensor) -> None:
        """Producer side: enqueue the copy of `block` (contiguous, <= slot_bytes) into the ring."""
        if self.arena.rank != self.src:
            raise RuntimeError("push is for the producer rank")
        if self.dead:
            raise N.NativeError("HbmRing: the ring failed earlier (timeout or oversized block) and cannot be used again")
        K.require_cuda(block)
        nbytes = block.numel() * block.element_size()
        a = self.arena
        N.check(N.kernels().dryad_ring_push(K.ptr(block), nbytes, a.peer_ptr(self.dst, self.ring_off), self._ctl(self.dst, 8),
                                            self._ctl(self.dst, 0), self._ctl(self.src, 1), self._ctl(self.src, 2),
                                            self.pushed, self.slots, self.slot_bytes, K.ptr(self.status), K.stream()), "ring_push")
        K._count()
        self.pushed += 1

    def pop(self, out: torch.Tensor) -> torch.Tensor:
        """Consumer side: enqueue the copy of the next block into `out` (capacity a multiple of 16
        bytes); returns the device word that will hold the block's byte length."""
        if self.arena.rank != self.dst:
            raise RuntimeError("pop is for the consumer rank")
        if self.dead:
            raise N.NativeError("HbmRing: the ring failed earlier (timeout or oversized block) and cannot be used again")
        K.require_cuda(out)
        a = self.arena
        cap = out.numel() * out.element_size()
        N.check(N.kernels().dryad_ring_pop(K.ptr(out), cap, K.ptr(self.out_len), a.peer_ptr(self.dst, self.ring_off),
                                           self._ctl(self.dst, 8), self._ctl(self.dst, 0), self._ctl(self.src, 1),
                                           self._ctl(self.dst, 3), self.popped, self.slots, self.slot_bytes,
                                           K.ptr(self.status), K.stream()), "ring_pop")
        K._count()
        self.popped += 1
        return self.out_len

    def check(self) -> None:
        st = int(self.status.item())
        if st:
            self.dead = True            # fatal: the per-block completion counters are out of step from here on
        if st == 2:
            raise N.NativeError("HbmRing: the peer did not move within the timeout")
        if st == 1:
            raise N.NativeError("HbmRing: block larger than the destination buffer")


def ring_send_batch(ring: HbmRing, batch, names) -> None:
    """Stream a ColumnBatch through the ring: one header block (row count), then every column in
    slot-sized chunks.  Names, dtypes and order are known to both ends from the plan's schema."""
    dev = ring.arena.device
    n = batch.n
    hdr = torch.tensor([n, 0], dtype=torch.int64, device=dev)
    ring.push(hdr)
    for k in names:
        col = batch.cols[k].contiguous().view(torch.uint8) if n else None
        if col is None:
            continue
        flat = col.reshape(-1)
        for a in range(0, flat.numel(), ring.slot_bytes):
            blk = flat[a:a + ring.slot_bytes]
            if blk.data_ptr() % 16:
                blk = blk.clone()
            ring.push(blk)


def ring_recv_batch(ring: HbmRing, names, dtypes) -> dict:
    """Inverse of ring_send_batch; returns dict name -> tensor."""
    dev = ring.arena.device
    hdr = torch.zeros(2, dtype=torch.int64, device=dev)
    ring.pop(hdr)
    n = int(hdr[0].item())                           # sizes the receives that follow
    out = {}
    for k in names:
        dt = dtypes[k]
        esz = torch.empty(0, dtype=dt).element_size()
        total = n * esz
        buf = torch.empty((total + 15) // 16 * 16 + ring.slot_bytes, dtype=torch.uint8, device=dev)
        for a in range(0, total, ring.slot_bytes):
            ring.pop(buf[a:a + ring.slot_bytes])
        out[k] = buf[:total].view(dt)
    ring.check()
    return out
