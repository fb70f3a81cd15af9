"""Distributed OrderBy over fixed-width rows: sample → separators → fused range shuffle → local
radix sort → tie repair → payload gather.

Mirrors the reference's OrderBy plan (LinqToDryad/DryadLinqQueryGen.cs:2362-2521): Phase1Sampling
per partition, a single RangeSampler producing `pcount-1` separators that are broadcast to every
distributor, RangePartition with CrossProduct edges into a Merge, then a per-partition Sort.
Here the sampler output crosses GPUs as small P2P copies, the RangePartition+Merge is the fused
shuffle kernel, and the sort is the one-sweep radix sort on (64-bit key prefix, row index) followed
by one payload gather, so each 100-byte row crosses NVLink once and HBM three times.
"""
from __future__ import annotations

from typing import Optional

import torch

from ..ops import core as K
from .shuffle import ColSpec, FusedShuffle
from .symm import SymmetricArena, tensor_from_ptr

SAMPLE_RATE = 0.001          # DryadLinqSampler.SAMPLE_RATE (DryadLinqSampler.cs:38)
MAX_SAMPLES = 1 << 20        # MAX_SECOND_PHASE_SAMPLES (DryadLinqSampler.cs:39)


class RowSorter:
    """Reusable distributed sorter for `[n, width]` uint8 rows keyed by `key_len` leading
    big-endian bytes at `key_off` (key_len >= 8; bytes beyond 8 are resolved by tie repair)."""

    def __init__(self, arena: SymmetricArena, max_rows: int, width: int, *, key_off: int = 0,
                 key_len: int = 10, slack: float = 1.12, sample_rate: float = SAMPLE_RATE,
                 descending: bool = False):
        if key_len < 8:
            raise ValueError("RowSorter needs at least 8 key bytes (pad shorter keys)")
        self.arena, self.width, self.key_off, self.key_len = arena, width, key_off, key_len
        self.descending = descending
        self.world, self.rank = arena.world, arena.rank
        self.max_rows = int(max_rows)
        dev = arena.device
        self.sample_rate = sample_rate
        self.sample_cap = max(1024, min(MAX_SAMPLES // self.world, int(max_rows * sample_rate * 2) + 1024))
        if self.world > 1:
            self.cap_rows = int(max_rows * slack) + 4096
            import os
            # DRYAD_SORT_EMIT_KEY=0: do not ship the 8-byte key prefix next to each row; the receiver
            # re-extracts it from the rows (7 % fewer NVLink bytes, one more pass over row heads)
            self.emit_key = os.environ.get("DRYAD_SORT_EMIT_KEY", "1") != "0"
            self.shuffle = FusedShuffle(arena, [ColSpec(width, torch.uint8, width)], self.cap_rows,
                                        emit_key=self.emit_key, max_rows_in=max_rows)
            self.all_samples, self.all_samples_off = arena.alloc((self.world, self.sample_cap), torch.int64)
            self.all_counts, self.all_counts_off = arena.alloc((self.world,), torch.int64)
            self._peer_sample_views = []
            self._peer_count_views = []
            for r in range(self.world):
                sp = tensor_from_ptr(arena.peer_ptr(r, self.all_samples_off + self.rank * self.sample_cap * 8),
                                     self.sample_cap * 8, dev).view(torch.int64)
                cp = tensor_from_ptr(arena.peer_ptr(r, self.all_counts_off + self.rank * 8), 8, dev).view(torch.int64)
                self._peer_sample_views.append(sp)
                self._peer_count_views.append(cp)
        else:
            self.cap_rows = self.max_rows
            self.shuffle = None
        n = self.cap_rows
        self.keys_a = torch.empty(n, dtype=torch.int64, device=dev)
        self.keys_b = torch.empty(n, dtype=torch.int64, device=dev)
        self.idx_a = torch.empty(n, dtype=torch.int32, device=dev)
        self.idx_b = torch.empty(n, dtype=torch.int32, device=dev)
        self.out = torch.empty((n, width), dtype=torch.uint8, device=dev)
        self.last_long_runs: Optional[torch.Tensor] = None
        self.profile = False                  # diagnostic: record per-phase device times in self.timing
        self.timing: dict = {}
        self._marks: list = []

    def _mark(self, name: str) -> None:
        if self.profile:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._marks.append((name, e))

    def _finish_marks(self) -> None:
        if self.profile and self._marks:
            torch.cuda.synchronize()
            self.timing = {b[0]: a[1].elapsed_time(b[1]) for a, b in zip(self._marks, self._marks[1:])}
            self._marks = []

    # -- separators ----------------------------------------------------------------------------
    def _separators(self, rows: torch.Tensor) -> torch.Tensor:
        samples, cnt = K.sample_rows(rows, self.key_off, seed=self.rank, rate=self.sample_rate,
                                     cap=self.sample_cap)
        cnt = torch.clamp(cnt, max=self.sample_cap)
        # publish my samples + count into every rank's table (P2P copies), then barrier
        for r in range(self.world):
            self._peer_sample_views[r].copy_(samples, non_blocking=True)
            self._peer_count_views[r].copy_(cnt, non_blocking=True)
        self.arena.barrier(slot=2)
        counts = self.all_counts.tolist()                      # host sync (tiny)
        parts = [self.all_samples[r, :int(counts[r])] for r in range(self.world)]
        allk = torch.cat(parts) if parts else samples[:0]
        total = allk.numel()
        skeys, _ = K.radix_sort_pairs(allk.contiguous(), None)   # ascending; the picker handles descending
        return K.select_separators(skeys, total, self.world, self.descending)

    # -- the sort ------------------------------------------------------------------------------
    WINDOW_BITS = 32          # radix passes cover this many key bits; ties are repaired from the rows
    WINDOW_MAX_RUN = 64

    @staticmethod
    def _window(lo: int, hi: int):
        """Bit window [b0, b1) of the 64-bit key prefix worth sorting when all keys lie in
        [lo, hi]: leading bits common to lo and hi are skipped."""
        top = max((lo ^ hi).bit_length(), RowSorter.WINDOW_BITS)
        return top - RowSorter.WINDOW_BITS, top

    def _local_sort(self, keys, m: int, src_rows: torch.Tensor, lo: int, hi: int):
        """Sort permutation of the first m rows of src_rows.  `keys`: the rows' 64-bit key prefixes
        (int64 bits) when the shuffle delivered them, else None.  Fast path: a 32-bit window of the
        prefix (leading bits shared by the whole partition are skipped) sorted as (u32, u32) pairs in
        4 radix passes + tie repair on the remaining key bytes; a partition with long runs of equal
        windows (not TeraSort-like) falls back to all 8 passes over the 64-bit prefix."""
        from .. import _native as N
        lib = N.kernels()
        b0, b1 = self._window(lo, hi)
        k32 = self.keys_a.view(torch.int32)[:m]
        if not self.descending:
            if keys is None:
                N.check(lib.dryad_extract_be_window(K.ptr(src_rows), m, self.width, self.key_off, b0, K.ptr(k32), K.stream()), "extract_be_window")
            else:
                N.check(lib.dryad_shift_u64_to_u32(K.ptr(keys), m, b0, K.ptr(k32), K.stream()), "shift_u64_to_u32")
            K._count()
            skeys, sidx = K.radix_sort_pairs(k32, self.idx_a[:m], iota=True, keys_alt=self.keys_b.view(torch.int32)[:m],
                                             vals_alt=self.idx_b[:m])
            first_byte = (64 - b0) // 8               # key bytes before this are equal inside a run
            runs = torch.zeros(1, dtype=torch.int32, device=src_rows.device)
            N.check(lib.dryad_fix_ties32(K.ptr(skeys), K.ptr(sidx), m, K.ptr(src_rows), self.width, self.key_off + first_byte,
                                         self.key_len - first_byte, self.WINDOW_MAX_RUN, 0, K.ptr(runs), K.stream()), "fix_ties32")
            K._count()
            if int(runs.item()) == 0:
                self.last_long_runs = runs
                return sidx
        # general path (descending order, or long runs of equal windows): full 64-bit prefix
        k64 = K.extract_be64(src_rows[:m], self.key_off, out=self.keys_a)[:m]
        if self.descending:
            k64 = ~k64
        skeys, sidx = K.radix_sort_pairs(k64, self.idx_a[:m], iota=True, keys_alt=self.keys_b[:m], vals_alt=self.idx_b[:m])
        if self.key_len > 8:
            self.last_long_runs = K.fix_ties(skeys, sidx, src_rows, self.key_off + 8, self.key_len - 8,
                                             descending=self.descending)
        return sidx

    def sort(self, rows: torch.Tensor):
        """rows: [n, width] uint8 on this rank.  Returns (sorted_rows_view, m): this rank's slice
        of the global order (rank r holds keys between separators r-1 and r)."""
        n = rows.shape[0]
        if n > self.max_rows:
            raise ValueError("RowSorter: more rows than the sorter was sized for")
        M64 = (1 << 64) - 1
        self._mark("start")
        if self.world > 1:
            seps = self._separators(rows)
            self._mark("sample+separators")
            mode = "range_desc" if self.descending else "range_asc"
            self.shuffle.run([rows], mode=mode, key_col=0, key_kind=1, key_off=self.key_off, separators=seps)
            self._mark("shuffle")
            sl = [x & M64 for x in seps.tolist()]              # host sync (tiny): this rank's key range
            m = self.shuffle.received_rows()[0]               # host sync: sizes the local sort
            self._mark("host sync")
            src_rows = self.shuffle.recv[0][0]
            keys = self.shuffle.recv_key[0][:m] if self.emit_key else None
            r = self.rank
            if self.descending:
                lo = sl[r] if r < len(sl) else 0
                hi = sl[r - 1] if r > 0 else M64
                lo, hi = (~hi) & M64, (~lo) & M64              # keys are complemented before sorting
            else:
                lo = sl[r - 1] if r > 0 else 0
                hi = sl[r] if r < len(sl) else M64
        else:
            m = n
            src_rows = rows
            keys = None
            lo, hi = 0, M64
        sidx = self._local_sort(keys, m, src_rows, lo, hi)
        self._mark("local sort + ties")
        out = self.out[:m]
        K.gather_rows(src_rows, sidx, out)
        self._mark("gather")
        self._finish_marks()
        return out, m
