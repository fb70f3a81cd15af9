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
        self._key_tmp = None
        if max_rows_in is not None:
            self._ensure_scratch(max_rows_in)

    def _ensure_scratch(self, n: int) -> torch.Tensor:
        need = N.kernels().dryad_shuffle_scratch_bytes(n, self.nparts)
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty(need, dtype=torch.uint8, device=self.arena.device)
        return self._scratch

    def _desc(self, srcs: Sequence[torch.Tensor], widths: Sequence[int], *, mode: str, key_col: int, key_kind: int,
              key_off: int, separators: Optional[torch.Tensor], epoch: int, dst_table: torch.Tensor,
              emit_key: bool) -> "N.ShuffleDesc":
        a = self.arena
        n = srcs[0].shape[0]
        K.require_cuda(*srcs)
        d = N.ShuffleDesc()
        d.ncols = len(srcs)
        d.nparts = self.nparts
        d.mode = K.PART[mode]
        d.key_col, d.key_kind, d.key_off = key_col, key_kind, key_off
        d.emit_key = 1 if emit_key else 0
        d.rank, d.world = a.rank, a.world
        d.n = n
        for c, (w, t) in enumerate(zip(widths, srcs)):
            if t.shape[0] != n:
                raise ValueError("shuffle columns must have equal row counts")
            row_bytes = t.element_size() * (t.shape[1] if t.dim() > 1 else 1)
            if row_bytes != w:
                raise ValueError(f"column {c}: row width {row_bytes} != spec {w}")
            d.src[c] = t.data_ptr()
            d.width[c] = w
        d.dst_table = dst_table.data_ptr()
        d.dst_key_table = self.dst_key_table.data_ptr() if emit_key else None
        if mode.startswith("range"):
            if separators is None or separators.numel() != self.nparts - 1:
                raise ValueError("range shuffle needs nparts-1 separators")
            d.separators = separators.data_ptr()
        d.capacity = self.capacity.data_ptr()
        d.counts_local = self.counts.data_ptr()
        d.counts_peers = self.counts_peers.data_ptr()
        d.flags_local = a.flags_local_ptr()
        d.flags_peers = a.flags_peers_table().data_ptr()
        d.epoch = epoch
        sc = self._ensure_scratch(n)
        d.scratch, d.scratch_bytes = sc.data_ptr(), sc.numel()
        d.status = self.status.data_ptr()
        return d

    def run(self, srcs: Sequence[torch.Tensor], *, mode: str, key_col: int = 0, key_kind: int = 0,
            key_off: int = 0, separators: Optional[torch.Tensor] = None) -> None:
        """Launch the fused shuffle on the current stream.  No host synchronisation.  Afterwards
        `self.counts[src_rank, part]` (device) holds the row counts and the receive buffers hold
        the rows, ordered by source rank then source order."""
        d = self._desc(srcs, [cs.width for cs in self.cols], mode=mode, key_col=key_col, key_kind=key_kind,
                       key_off=key_off, separators=separators, epoch=self.arena.next_epoch(),
                       dst_table=self.dst_table, emit_key=self.emit_key)
        N.check(N.kernels().dryad_shuffle_run(C.byref(d), K.stream()), "shuffle_run")
        K._count(4)

    def run_passes(self, rows: torch.Tensor, *, mode: str, key_off: int = 0,
                   separators: Optional[torch.Tensor] = None, side: "torch.cuda.Stream" = None,
                   payload_ctas: int = 0) -> list:
        """Key stream first, then the wide rows one key sub-range per pass (single payload column,
        emit_key=True buffers, range mode with parts_per_rank = R sub-ranges per destination).

        Current stream: count + publish (the count pass also writes the rows' 64-bit key prefixes as
        a compact column), then the 8-byte keys of ALL partitions are scattered into `recv_key` and
        barrier slot 1 closes the key stream.  `side` stream: R passes of the TMA scatter over the
        rows with the SAME offsets (slot i of recv_key[j] is row i of recv[0][j]); pass j moves only
        the partitions of sub-range j and is closed by its own barrier (slot 4).  Returns one event
        per pass: after event j every destination's sub-range-j buffer is complete, so the consumer
        can sort keys while the link-bound row transfer runs and finish sub-range j while pass j+1
        is still in flight.  Re-reading the rows R times is free where the link is the limit."""
        if not self.emit_key or len(self.cols) != 1:
            raise ValueError("run_passes needs one payload column and emit_key=True")
        n = rows.shape[0]
        a = self.arena
        if self._key_tmp is None or self._key_tmp.numel() < max(n, 1):
            self._key_tmp = torch.empty(max(n, 1), dtype=torch.int64, device=a.device)
        epoch = a.next_epoch()
        lib = N.kernels()
        cur = torch.cuda.current_stream()
        w = self.cols[0].width
        d = self._desc([rows], [w], mode=mode, key_col=0, key_kind=1, key_off=key_off, separators=separators,
                       epoch=epoch, dst_table=self.dst_table, emit_key=False)
        d.key_out = self._key_tmp.data_ptr()
        d.phases = N.SHUFFLE_PLAN
        N.check(lib.dryad_shuffle_run(C.byref(d), K.stream()), "shuffle_plan")
        planned = torch.cuda.Event()
        planned.record(cur)
        dk = self._desc([self._key_tmp[:n]], [8], mode=mode, key_col=0, key_kind=0, key_off=0, separators=separators,
                        epoch=epoch, dst_table=self.dst_key_table, emit_key=False)
        dk.phases = N.SHUFFLE_SCATTER | N.SHUFFLE_BARRIER
        dk.barrier_slot = 1
        N.check(lib.dryad_shuffle_run(C.byref(dk), K.stream()), "shuffle_keys")
        side.wait_event(planned)
        d.key_out = None
        d.phases = N.SHUFFLE_SCATTER | N.SHUFFLE_BARRIER
        d.barrier_slot = 4
        d.max_ctas = int(payload_ctas)
        events = []
        for j in range(self.ppr):
            d.filter_mod, d.filter_rem = (self.ppr, j) if self.ppr > 1 else (0, 0)
            d.barrier_epoch = a.next_epoch()
            N.check(lib.dryad_shuffle_run(C.byref(d), side.cuda_stream), "shuffle_payload")
            ev = torch.cuda.Event()
            ev.record(side)
            events.append(ev)
        K._count(4 + 2 * self.ppr)
        return events

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
