"""Distributed OrderBy over fixed-width rows: sample → separators → range exchange → local radix
sort → tie repair → payload gather.

Mirrors the reference's OrderBy plan (LinqToDryad/DryadLinqQueryGen.cs:2362-2521): Phase1Sampling
per partition, a single RangeSampler producing `pcount-1` separators that are broadcast to every
distributor, RangePartition with CrossProduct edges into a Merge, then a per-partition Sort.
Here the sampler output crosses GPUs as small P2P copies, the RangePartition+Merge is the sort
exchange of csrc/kernels/sortx.cu (ascending order; keys first, then the rows one key sub-range per
round, so the receiver sorts the keys and gathers finished sub-ranges while later rounds are still
on NVLink) or the fused shuffle kernel (descending order), and the sort is the one-sweep radix sort
on (key window, slot) followed by a payload gather: each 100-byte row crosses NVLink once.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import _native as N
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
                 descending: bool = False, passes: Optional[int] = None):
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
            import os
            # every destination's key range can be cut into R sub-ranges with their own receive
            # buffers (FusedShuffle parts_per_rank): the rows then travel one sub-range per pass and a
            # sub-range is gathered while later passes are still on NVLink.  Measured on 2 and 8
            # B200s each extra pass costs more SM time re-reading the rows than the overlap returns
            # (profiles/README.md), so the default is one pass.
            R = passes if passes is not None else int(os.environ.get("DRYAD_SORT_PASSES", "1"))
            self.R = max(1, min(16, R, 256 // self.world))
            # keys first / rows behind on a side stream (ascending order); 0 = one fused shuffle
            self.split = os.environ.get("DRYAD_SORT_SPLIT", "1") != "0" and not descending
            # the row transfer is link-bound: a capped number of persistent CTAs leaves SM resources
            # to the key sort and the gathers that run concurrently on the main stream
            self.payload_ctas = int(os.environ.get("DRYAD_SORT_PAYLOAD_CTAS", "148"))
            self.cap_part = int(max_rows * slack / self.R) + 4096
            self.cap_rows = self.cap_part * self.R
            self.emit_key = True
            self.shuffle = FusedShuffle(arena, [ColSpec(width, torch.uint8, width)], self.cap_part,
                                        parts_per_rank=self.R, emit_key=True, max_rows_in=max_rows)
            self.regrown = 0
            self.side = torch.cuda.Stream(device=dev) if self.split else None
            self._peek_buf = torch.empty(1024, dtype=torch.int64).pin_memory()
            self._peek_ev = torch.cuda.Event()
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
            self._samples = torch.empty(self.sample_cap, dtype=torch.int64, device=dev)
            self._sample_cnt = torch.zeros(1, dtype=torch.int64, device=dev)
            self._sorted_alt = torch.empty(self.world * self.sample_cap, dtype=torch.int64, device=dev)
            self._seps = torch.empty(self.world * self.R - 1, dtype=torch.int64, device=dev)
            self._tie_flags = torch.zeros((self.R, 2), dtype=torch.int32, device=dev)   # per sub-range: [long runs, unresolved]
        else:
            self.R = 1
            self.cap_part = self.cap_rows = self.max_rows
            self.shuffle = None
            self.split = False
        self._alloc_work()
        self._runs = torch.zeros(1, dtype=torch.int32, device=dev)
        self.last_long_runs: Optional[torch.Tensor] = None
        self.profile = False                  # diagnostic: record per-phase device times in self.timing
        self.timing: dict = {}
        self._marks: list = []

    def _alloc_work(self) -> None:
        n, dev = self.cap_rows, self.arena.device
        self.keys_a = torch.empty(n, dtype=torch.int64, device=dev)
        self.keys_b = torch.empty(n, dtype=torch.int64, device=dev)
        self.idx_a = torch.empty(n, dtype=torch.int32, device=dev)
        self.idx_b = torch.empty(n, dtype=torch.int32, device=dev)
        self.out = torch.empty((n, self.width), dtype=torch.uint8, device=dev)

    def _grow(self, need_rows: int) -> None:
        """Skew: one key range received more rows than its buffer holds (duplicate-heavy keys put a
        whole run of equal keys into ONE partition, whatever the separators).  The exact per-partition
        totals are known to every rank before any row moves (the shuffle's published count matrix),
        so all ranks take the same decision: re-create the receive side with room for the largest
        partition and run the exchange again.  Reference: consumers sized from the measured data
        (GraphManager/stagemanager/DrDynamicDistributor.h:25-124); file channels absorb any skew."""
        self.cap_part = int(need_rows * 1.05) + 4096
        self.cap_rows = self.cap_part * self.R
        self.keys_a = self.keys_b = self.idx_a = self.idx_b = self.out = None
        self.shuffle = FusedShuffle(self.arena, [ColSpec(self.width, torch.uint8, self.width)], self.cap_part,
                                    parts_per_rank=self.R, emit_key=True, max_rows_in=self.max_rows)
        self._alloc_work()
        self.regrown += 1

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
    def _peek(self, t: torch.Tensor) -> list:
        """Read a small device tensor on the host without draining the stream behind it: async copy
        into pinned memory, then wait for THAT copy only (work queued later, or on the side
        stream, keeps running)."""
        k = t.numel()
        self._peek_buf[:k].copy_(t, non_blocking=True)
        self._peek_ev.record()
        self._peek_ev.synchronize()
        return self._peek_buf[:k].tolist()

    def _separators(self, rows: torch.Tensor) -> torch.Tensor:
        """Phase1Sampling + RangeSampler (DryadLinqSampler.cs:42-106, 139-246) with no host round
        trip: every rank writes a fixed-size sample slab (unused slots = ~0, which sort last) and its
        count into every rank's table; each rank sorts the whole table and picks the world*R - 1
        separators from the first sum(counts) entries."""
        lib = N.kernels()
        self._samples.fill_(-1)
        self._sample_cnt.zero_()
        N.check(lib.dryad_sample_rows(K.ptr(rows), rows.shape[1], self.key_off, rows.shape[0], self.rank,
                                      self.sample_rate, K.ptr(self._samples), K.ptr(self._sample_cnt),
                                      self.sample_cap, K.stream()), "sample_rows")
        K._count()
        for r in range(self.world):
            self._peer_sample_views[r].copy_(self._samples, non_blocking=True)
            self._peer_count_views[r].copy_(self._sample_cnt, non_blocking=True)
        self.arena.barrier(slot=2)
        allk = self.all_samples.view(-1)
        skeys, _ = K.radix_sort_pairs(allk, None, keys_alt=self._sorted_alt)   # ascending; the picker handles descending
        N.check(lib.dryad_select_separators_dev(K.ptr(skeys), K.ptr(self.all_counts), self.world, self.sample_cap,
                                                self.world * self.R, int(self.descending), K.ptr(self._seps), K.stream()),
                "select_separators_dev")
        K._count()
        return self._seps

    # -- the sort ------------------------------------------------------------------------------
    WINDOW_BITS = 32          # radix passes cover this many key bits; ties are repaired afterwards
    WINDOW_MAX_RUN = 64

    @staticmethod
    def _window(lo: int, hi: int):
        """Bit window [b0, b1) of the 64-bit key prefix worth sorting when all keys lie in
        [lo, hi]: leading bits common to lo and hi are skipped."""
        top = max((lo ^ hi).bit_length(), RowSorter.WINDOW_BITS)
        return top - RowSorter.WINDOW_BITS, top

    def _local_sort(self, keys, m: int, src_rows: torch.Tensor, lo: int, hi: int, o: int = 0):
        """Sort permutation of the first m rows of src_rows, computed in the work buffers at offset
        `o`.  `keys`: the rows' 64-bit key prefixes (int64 bits) when the shuffle delivered them,
        else None.  Fast path: a 32-bit window of the prefix (leading bits shared by the whole
        partition are skipped) sorted as (u32, u32) pairs in 4 radix passes + tie repair on the
        remaining key bytes read from the rows; a partition with long runs of equal windows (not
        TeraSort-like) falls back to all 8 passes over the 64-bit prefix."""
        lib = N.kernels()
        b0, b1 = self._window(lo, hi)
        k32a, k32b = self.keys_a.view(torch.int32), self.keys_b.view(torch.int32)
        if not self.descending:
            k32 = k32a[o:o + m]
            if keys is None:
                N.check(lib.dryad_extract_be_window(K.ptr(src_rows), m, self.width, self.key_off, b0, K.ptr(k32), K.stream()), "extract_be_window")
            else:
                N.check(lib.dryad_shift_u64_to_u32(K.ptr(keys), m, b0, K.ptr(k32), K.stream()), "shift_u64_to_u32")
            K._count()
            skeys, sidx = K.radix_sort_pairs(k32, self.idx_a[o:o + m], iota=True, keys_alt=k32b[o:o + m],
                                             vals_alt=self.idx_b[o:o + m])
            first_byte = (64 - b0) // 8               # key bytes before this are equal inside a run
            runs = self._runs
            runs.zero_()
            N.check(lib.dryad_fix_ties32(K.ptr(skeys), K.ptr(sidx), m, K.ptr(src_rows), self.width, self.key_off + first_byte,
                                         self.key_len - first_byte, self.WINDOW_MAX_RUN, 0, K.ptr(runs), K.stream()), "fix_ties32")
            K._count()
            if int(runs.item()) == 0:
                self.last_long_runs = runs
                return sidx
        # general path (descending order, or long runs of equal windows): full 64-bit prefix
        k64 = K.extract_be64(src_rows[:m], self.key_off, out=self.keys_a[o:o + m])
        if self.descending:
            k64 = ~k64
        skeys, sidx = K.radix_sort_pairs(k64, self.idx_a[o:o + m], iota=True, keys_alt=self.keys_b[o:o + m],
                                         vals_alt=self.idx_b[o:o + m])
        if self.key_len > 8:
            self.last_long_runs = K.fix_ties(skeys, sidx, src_rows, self.key_off + 8, self.key_len - 8,
                                             descending=self.descending)
        return sidx

    def _sort_keys(self, keys, m: int, lo: int, hi: int, o: int, flags: torch.Tensor):
        """Sort permutation from the received key prefixes ALONE (32-bit window + key-only tie
        repair), so it can run before the rows have arrived.  flags[0] / flags[1] count what the
        keys cannot decide (long runs of equal windows / equal 64-bit prefixes under a wider key)."""
        lib = N.kernels()
        b0, _ = self._window(lo, hi)
        k32 = self.keys_a.view(torch.int32)[o:o + m]
        N.check(lib.dryad_shift_u64_to_u32(K.ptr(keys), m, b0, K.ptr(k32), K.stream()), "shift_u64_to_u32")
        skeys, sidx = K.radix_sort_pairs(k32, self.idx_a[o:o + m], iota=True,
                                         keys_alt=self.keys_b.view(torch.int32)[o:o + m], vals_alt=self.idx_b[o:o + m])
        N.check(lib.dryad_fix_ties32_keys(K.ptr(skeys), K.ptr(sidx), m, K.ptr(keys), self.WINDOW_MAX_RUN,
                                          1 if self.key_len > 8 else 0, flags.data_ptr(), flags.data_ptr() + 4, K.stream()),
                "fix_ties32_keys")
        K._count(2)
        return sidx

    def sort(self, rows: torch.Tensor):
        """rows: [n, width] uint8 on this rank.  Returns (sorted_rows_view, m): this rank's slice
        of the global order (rank r holds keys between separators r*R-1 and (r+1)*R-1)."""
        n = rows.shape[0]
        if n > self.max_rows:
            raise ValueError("RowSorter: more rows than the sorter was sized for")
        M64 = (1 << 64) - 1
        self._mark("start")
        if self.world == 1:
            sidx = self._local_sort(None, n, rows, 0, M64)
            self._mark("local sort + ties")
            out = self.out[:n]
            K.gather_rows(rows, sidx, out)
            self._mark("gather")
            self._finish_marks()
            return out, n

        W, r, R = self.world, self.rank, self.R
        sh = self.shuffle
        main = torch.cuda.current_stream()
        seps = self._separators(rows)
        self._mark("sample+separators")
        mode = "range_desc" if self.descending else "range_asc"
        events = None
        if self.split:
            events = sh.run_passes(rows, mode=mode, key_off=self.key_off, separators=seps, side=self.side,
                                   payload_ctas=self.payload_ctas)
            self._mark("plan + key stream")
        else:
            sh.run([rows], mode=mode, key_col=0, key_kind=1, key_off=self.key_off, separators=seps)
            self._mark("shuffle")
        # ONE small read-back: status, rows received per (source, my sub-range), the separators
        info = self._peek(torch.cat([sh.status.to(torch.int64), sh.counts[:, r * R:(r + 1) * R].reshape(-1), seps]))
        st = int(info[0])
        if st == 1:
            # every rank sees the same count matrix, so every rank grows and re-runs together
            if events is not None:
                main.wait_event(events[-1])
            need = int(sh.counts.sum(dim=0).max().item())
            if need <= self.cap_part or self.regrown >= 4:
                raise N.NativeError("shuffle: a destination partition overflowed its receive buffer")
            sh.status.zero_()
            self._marks = []
            self._grow(need)
            return self.sort(rows)
        if st == 2:
            raise N.NativeError("shuffle: peer did not signal within the timeout")
        cnt = info[1:1 + W * R]
        mj = [int(sum(cnt[s * R + j] for s in range(W))) for j in range(R)]
        sl = [x & M64 for x in info[1 + W * R:]]
        self._mark("host sync")
        base = [0]
        for j in range(R):
            base.append(base[-1] + mj[j])
        m = base[R]

        def key_range(j: int):
            p = r * R + j                               # my j-th partition in the global order
            if self.descending:
                lo = sl[p] if p < len(sl) else 0
                hi = sl[p - 1] if p > 0 else M64
                return (~hi) & M64, (~lo) & M64          # keys are complemented before sorting
            return (sl[p - 1] if p > 0 else 0), (sl[p] if p < len(sl) else M64)

        if events is not None:
            # all key sorts first (they need the keys only), then each sub-range's gather as soon as
            # its pass has landed
            self._tie_flags.zero_()
            perms = []
            for j in range(R):
                lo, hi = key_range(j)
                perms.append(self._sort_keys(sh.recv_key[j][:mj[j]], mj[j], lo, hi, j * self.cap_part, self._tie_flags[j])
                             if mj[j] else None)
            self._mark("key sorts")
            bad = self._peek(self._tie_flags.reshape(-1))
            for j in range(R):
                main.wait_event(events[j])
                if not mj[j]:
                    continue
                src = sh.recv[0][j]
                sidx = perms[j]
                if bad[2 * j] or bad[2 * j + 1]:            # the keys alone could not decide: row-based path
                    lo, hi = key_range(j)
                    sidx = self._local_sort(sh.recv_key[j][:mj[j]], mj[j], src, lo, hi, j * self.cap_part)
                K.gather_rows(src, sidx, self.out[base[j]:base[j + 1]])
            self._mark("row passes + gathers")
        else:
            for j in range(R):
                if not mj[j]:
                    continue
                lo, hi = key_range(j)
                src = sh.recv[0][j]
                sidx = self._local_sort(sh.recv_key[j][:mj[j]], mj[j], src, lo, hi, j * self.cap_part)
                K.gather_rows(src, sidx, self.out[base[j]:base[j + 1]])
            self._mark("local sorts + gathers")
        self._finish_marks()
        return self.out[:m], m
