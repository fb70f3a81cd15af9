"""Fused partition + all-to-all over peer memory (K1/K2 + C1).

One `FusedShuffle` object owns the receive buffers (in the symmetric arena), the count matrix and
the pointer tables for a fixed schema, and can be run repeatedly.  Partition p of `nparts` lives on
rank `p // parts_per_rank` — contiguous, so a range partition leaves partitions globally ordered by
(rank, local index), which is what OrderBy relies on (LinqToDryad/DryadLinqQueryGen.cs:2476-2521).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from .. import _native as N
from ..ops import core as K
from .symm import SymmetricArena


@dataclass
class ColSpec:
    width: int           # bytes per row (multiple of 4)
    dtype: torch.dtype   # element dtype of the receive buffer view
    inner: int = 1       # elements per row


class FusedShuffle:
    def __init__(self, arena: SymmetricArena, cols: Sequence[ColSpec], cap_rows: int, *,
                 parts_per_rank: int = 1, emit_key: bool = False, max_rows_in: Optional[int] = None):
        self.arena = arena
        self.cols = list(cols)
        self.ppr = parts_per_rank
        self.nparts = arena.world * parts_per_rank
        if self.nparts > 256:
            raise ValueError("at most 256 destination partitions")
        self.cap_rows = int(cap_rows)
        self.emit_key = emit_key
        dev = arena.device
        W = arena.world
        # receive buffers: per column, per local partition
        self.recv: List[List[torch.Tensor]] = []
        self.recv_off: List[List[int]] = []
        for cs in self.cols:
            bufs, offs = [], []
            for _ in range(self.ppr):
                off = arena.alloc_bytes(max(self.cap_rows, 1) * cs.width)
                offs.append(off)
                shape = (self.cap_rows, cs.inner) if cs.inner > 1 else (self.cap_rows,)
                bufs.append(arena.view(off, shape, cs.dtype))
            self.recv.append(bufs)
            self.recv_off.append(offs)
        self.recv_key: List[torch.Tensor] = []
        self.recv_key_off: List[int] = []
        if emit_key:
            for _ in range(self.ppr):
                off = arena.alloc_bytes(max(self.cap_rows, 1) * 8)
                self.recv_key_off.append(off)
                self.recv_key.append(arena.view(off, (self.cap_rows,), torch.int64))
        self.counts, self.counts_off = arena.alloc((W, self.nparts), torch.int64)
        arena.check_symmetric("FusedShuffle receive buffers")
        # pointer tables (device)
        dst = []
        for c in range(len(self.cols)):
            for p in range(self.nparts):
                dst.append(arena.peer_ptr(p // self.ppr, self.recv_off[c][p % self.ppr]))
        self.dst_table = torch.tensor(dst, dtype=torch.int64).to(dev)
        if emit_key:
            self.dst_key_table = torch.tensor(
                [arena.peer_ptr(p // self.ppr, self.recv_key_off[p % self.ppr]) for p in range(self.nparts)],
                dtype=torch.int64).to(dev)
        else:
            self.dst_key_table = None
        self.counts_peers = arena.peer_ptr_table(self.counts_off)
        self.capacity = torch.full((self.nparts,), self.cap_rows, dtype=torch.int64, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self._scratch = None
        if max_rows_in is not None:
            self._ensure_scratch(max_rows_in)

    def _ensure_scratch(self, n: int) -> torch.Tensor:
        need = N.kernels().dryad_shuffle_scratch_bytes(n, self.nparts)
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty(need, dtype=torch.uint8, device=self.arena.device)
        return self._scratch

    def run(self, srcs: Sequence[torch.Tensor], *, mode: str, key_col: int = 0, key_kind: int = 0,
            key_off: int = 0, separators: Optional[torch.Tensor] = None) -> None:
        """Launch the fused shuffle on the current stream.  No host synchronisation.  Afterwards
        `self.counts[src_rank, part]` (device) holds the row counts and the receive buffers hold
        the rows, ordered by source rank then source order."""
        a = self.arena
        n = srcs[0].shape[0]
        K.require_cuda(*srcs)
        d = N.ShuffleDesc()
        d.ncols = len(self.cols)
        d.nparts = self.nparts
        d.mode = K.PART[mode]
        d.key_col, d.key_kind, d.key_off = key_col, key_kind, key_off
        d.emit_key = 1 if self.emit_key else 0
        d.rank, d.world = a.rank, a.world
        d.n = n
        for c, (cs, t) in enumerate(zip(self.cols, srcs)):
            if t.shape[0] != n:
                raise ValueError("shuffle columns must have equal row counts")
            row_bytes = t.element_size() * (t.shape[1] if t.dim() > 1 else 1)
            if row_bytes != cs.width:
                raise ValueError(f"column {c}: row width {row_bytes} != spec {cs.width}")
            d.src[c] = t.data_ptr()
            d.width[c] = cs.width
        d.dst_table = self.dst_table.data_ptr()
        d.dst_key_table = self.dst_key_table.data_ptr() if self.emit_key else None
        if mode.startswith("range"):
            if separators is None or separators.numel() != self.nparts - 1:
                raise ValueError("range shuffle needs nparts-1 separators")
            d.separators = separators.data_ptr()
        d.capacity = self.capacity.data_ptr()
        d.counts_local = self.counts.data_ptr()
        d.counts_peers = self.counts_peers.data_ptr()
        d.flags_local = a.flags_local_ptr()
        d.flags_peers = a.flags_peers_table().data_ptr()
        d.epoch = a.next_epoch()
        sc = self._ensure_scratch(n)
        d.scratch, d.scratch_bytes = sc.data_ptr(), sc.numel()
        d.status = self.status.data_ptr()
        N.check(N.kernels().dryad_shuffle_run(C.byref(d), K.stream()), "shuffle_run")
        K._count(4)

    def received_rows(self) -> List[int]:
        """Rows received per local partition (host sync)."""
        st = int(self.status.item())
        if st == 1:
            raise N.NativeError("shuffle: a destination partition overflowed its receive buffer")
        if st == 2:
            raise N.NativeError("shuffle: peer did not signal within the timeout")
        tot = self.counts.sum(dim=0).tolist()
        r = self.arena.rank
        return [int(tot[r * self.ppr + j]) for j in range(self.ppr)]

    # ---- debug protocol checker ---------------------------------------------------------------
    def poison(self) -> None:
        """Fill the receive buffers with 0xA5 so that rows a broken handshake failed to deliver
        (or delivered to the wrong place) cannot masquerade as data."""
        for bufs in self.recv:
            for b in bufs:
                b.view(torch.uint8).fill_(0xA5)
        for b in self.recv_key:
            b.view(torch.uint8).fill_(0xA5)

    @staticmethod
    def _checksum(t: torch.Tensor, rows: int) -> torch.Tensor:
        if rows == 0:
            return torch.zeros(2, dtype=torch.int64, device=t.device)
        v = t[:rows].contiguous().view(torch.int32).to(torch.int64)
        w = (torch.arange(v.shape[-1] if v.dim() > 1 else 1, device=t.device, dtype=torch.int64) * 2654435761 + 1) & 0xFFFFFFFF
        v = v.view(rows, -1)
        return torch.stack([v.sum(), (v * w[: v.shape[1]]).sum()])          # wraps modulo 2^64

    def verify(self, srcs: Sequence[torch.Tensor]) -> None:
        """After run(): every row that was sent arrived exactly once somewhere — global row count and
        two position-weighted checksums per column over ALL ranks' received rows equal those of all
        sent rows (order-independent, so it holds for any correct partition function).  Collective
        over the process group; raises on mismatch.  Debug aid (`DRYAD_SHUFFLE_VERIFY=1` in tests),
        not part of the hot path."""
        import torch.distributed as dist
        got = self.received_rows()
        dev = self.arena.device
        sent = [torch.tensor([srcs[0].shape[0], 0], dtype=torch.int64, device=dev)]
        recv = [torch.tensor([sum(got), 0], dtype=torch.int64, device=dev)]
        for c, t in enumerate(srcs):
            sent.append(self._checksum(t, t.shape[0]))
            acc = torch.zeros(2, dtype=torch.int64, device=dev)
            for j in range(self.ppr):
                acc = acc + self._checksum(self.recv[c][j], got[j])
            recv.append(acc)
        a, b = torch.stack(sent), torch.stack(recv)
        if self.arena.world > 1:
            dist.all_reduce(a)
            dist.all_reduce(b)
        if not torch.equal(a, b):
            raise N.NativeError(f"shuffle verification failed: sent {a.tolist()} received {b.tolist()}")
