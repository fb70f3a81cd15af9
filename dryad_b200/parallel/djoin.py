"""Distributed joins.

`BroadcastJoin` (K8 + C4): the small side is replicated to every GPU — one NVLS multicast store per
16 bytes when the box has NVSwitch multicast (parallel/nvls.py), else P2P stores to every peer — the
hash table is built locally over the replica, then every rank probes its own fact partition: the
large side never moves.  The reference expresses this as Tee + CrossProduct + Merge(parCount) feeding
a HashJoin (LinqToDryad/DryadLinqQueryGen.cs:2957-3001); the planner rule that picks it by size is
QueryGen._visit_join.

`shuffle_hash_join`: both sides hash-partitioned by key with the fused shuffle, then a local hash
join (the reference's default plan, DryadLinqQueryGen.cs:1483-1608).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Sequence

import torch

from .. import _native as N
from ..ops import core as K
from .shuffle import ColSpec, FusedShuffle
from .symm import SymmetricArena


class BroadcastJoin:
    def __init__(self, arena: SymmetricArena, dim_rows_total: int, ncols: int, load: float = 0.5, use_nvls: bool = True):
        self.arena, self.world, self.rank = arena, arena.world, arena.rank
        self.nrows, self.ncols = int(dim_rows_total), ncols
        self.cap = K._next_pow2(int(dim_rows_total / load) + 64)
        dev = arena.device
        self.keys, ko = arena.alloc((self.cap,), torch.int64)
        self.head, ho = arena.alloc((self.cap,), torch.int32)
        self.next, no = arena.alloc((max(self.nrows, 1),), torch.int32)
        # the replicated small side lives in an NVLS buffer when the box has multicast: one store
        # per 16 bytes reaches every GPU (K17 nvls_broadcast_copy) instead of `world` P2P stores
        from .nvls import NvlsBuffer
        nr = max(self.nrows, 1)
        self.nvls = NvlsBuffer.create((ncols + 1) * (nr * 8 + 256) + 4096, dev, self.rank, self.world) if use_nvls else None
        self.cols, offs = [], []
        if self.nvls is not None:
            self.rkeys, self._nv_key_off = self.nvls.alloc((nr,), torch.int64)
            self._nv_col_off = []
            for _ in range(ncols):
                t, o = self.nvls.alloc((nr,), torch.int64)
                self.cols.append(t)
                self._nv_col_off.append(o)
            rko = 0
            offs = [0] * ncols
        else:
            self.rkeys, rko = arena.alloc((nr,), torch.int64)     # replicated key column
            for _ in range(ncols):
                t, o = arena.alloc((nr,), torch.int64)
                self.cols.append(t)
                offs.append(o)
        arena.check_symmetric("BroadcastJoin tables")
        W = self.world
        self._pk = (C.c_void_p * W)(*[arena.peer_ptr(r, ko) for r in range(W)])
        self._ph = (C.c_void_p * W)(*[arena.peer_ptr(r, ho) for r in range(W)])
        self._pn = (C.c_void_p * W)(*[arena.peer_ptr(r, no) for r in range(W)])
        self._pc = (C.c_void_p * max(W * ncols, 1))(*[arena.peer_ptr(r, offs[c]) for r in range(W) for c in range(ncols)])
        self._prepl = (C.c_void_p * (W * (ncols + 1)))(*[arena.peer_ptr(r, o) for r in range(W) for o in [rko] + offs])
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)

    def build(self, keys_local: torch.Tensor, cols_local: Sequence[torch.Tensor], row_base: int) -> None:
        """Replicate this rank's slice of the small side (key + payload columns) to every rank with
        P2P stores, then build the hash table over the full replica LOCALLY (local atomics are far
        cheaper than remote ones).  `row_base`: global offset of this rank's slice."""
        lib = N.kernels()
        if self.world > 1:
            self.arena.barrier(slot=2)             # nobody still probes the previous replica
        nl = keys_local.numel()
        if self.nvls is not None and row_base % 2 == 0 and nl % 2 == 0:
            self.nvls.broadcast(self._nv_key_off + row_base * 8, keys_local.contiguous())
            for c, o in zip(cols_local, self._nv_col_off):
                self.nvls.broadcast(o + row_base * 8, c.contiguous())
        elif self.nvls is not None:
            # odd slice boundaries are not 16-byte aligned: plain stores into every copy
            from .symm import tensor_from_ptr
            dev = keys_local.device
            for r in range(self.world):
                for t, o in [(keys_local, self._nv_key_off)] + list(zip(cols_local, self._nv_col_off)):
                    tensor_from_ptr(self.nvls.peer_ptrs[r] + o + row_base * 8, nl * 8, dev).view(torch.int64).copy_(t, non_blocking=True)
        else:
            src = (C.c_void_p * (self.ncols + 1))(keys_local.data_ptr(), *[c.data_ptr() for c in cols_local])
            N.check(lib.dryad_replicate_slice(self.world, self.ncols + 1, src, self._prepl, nl, row_base,
                                              K.stream()), "replicate_slice")
            K._count()
        if self.world > 1:
            self.arena.barrier(slot=3)             # all slices have landed everywhere
        N.check(lib.dryad_join_build(K.ptr(self.keys), K.ptr(self.head), self.cap, K.ptr(self.rkeys), K.ptr(self.next),
                                     self.nrows, K.ptr(self.status), K.stream()), "join_build")
        K._count(2)

    def slot_payload(self, col: torch.Tensor) -> torch.Tensor:
        """`col` (one value per small-side row) permuted into TABLE-SLOT order: out[s] = col[head[s]]
        for occupied slots.  A fused probe that returns the slot (`expr.probe_slot`) then reads the
        payload with ONE more memory request per row instead of two (head[slot], then col[row])."""
        idx = self.head.clamp(min=0)
        return K.gather_rows(col.contiguous(), idx.contiguous())

    def build_remote_atomics(self, keys_local: torch.Tensor, cols_local: Sequence[torch.Tensor], row_base: int) -> None:
        """Alternative single-kernel form: insert into every rank's table with system-scope atomics
        over NVLink (kept for comparison; slower at 8 GPUs than replicate + local build)."""
        lib = N.kernels()
        if self.nvls is not None:
            raise RuntimeError("build_remote_atomics needs the P2P replica layout: construct with use_nvls=False")
        N.check(lib.dryad_join_table_init(K.ptr(self.keys), K.ptr(self.head), self.cap, K.stream()), "join_table_init")
        K._count()
        if self.world > 1:
            self.arena.barrier(slot=2)             # every table is reset before any rank inserts
        src = (C.c_void_p * max(self.ncols, 1))(*[c.data_ptr() for c in cols_local])
        N.check(lib.dryad_join_build_bcast(K.ptr(keys_local), keys_local.numel(), row_base, self.world, self._pk,
                                           self._ph, self._pn, self.cap, self.ncols, src, self._pc,
                                           K.ptr(self.status), K.stream()), "join_build_bcast")
        K._count()
        if self.world > 1:
            self.arena.barrier(slot=3)             # all slices have landed everywhere

    def probe(self, probe_keys: torch.Tensor):
        """(probe_idx, dim_row) int32 pairs for this rank's probe keys against the full small side."""
        n = probe_keys.numel()
        dev = probe_keys.device
        lib = N.kernels()
        counts = torch.empty(n, dtype=torch.int64, device=dev)
        N.check(lib.dryad_join_probe(K.ptr(self.keys), K.ptr(self.head), self.cap, K.ptr(self.next), K.ptr(probe_keys),
                                     n, 0, K.ptr(counts), None, None, K.stream()), "join_probe")
        offs, tot = K.exclusive_scan_i64(counts)
        total = int(tot.item()) if n else 0
        op = torch.empty(total, dtype=torch.int32, device=dev)
        ob = torch.empty(total, dtype=torch.int32, device=dev)
        if total:
            N.check(lib.dryad_join_probe(K.ptr(self.keys), K.ptr(self.head), self.cap, K.ptr(self.next),
                                         K.ptr(probe_keys), n, 1, K.ptr(offs), K.ptr(op), K.ptr(ob), K.stream()), "join_probe")
        K._count(2)
        return op, ob


    def probe_unique(self, probe_keys: torch.Tensor):
        """For a small side with unique keys: (dim_row int32 [n], matched uint8 [n]) in one pass."""
        n = probe_keys.numel()
        rows = torch.empty(n, dtype=torch.int32, device=probe_keys.device)
        matched = torch.empty(n, dtype=torch.uint8, device=probe_keys.device)
        N.check(N.kernels().dryad_join_probe_unique(K.ptr(self.keys), K.ptr(self.head), self.cap, K.ptr(probe_keys), n,
                                                    K.ptr(rows), K.ptr(matched), K.stream()), "join_probe_unique")
        K._count()
        return rows, matched


class ShuffleHashJoin:
    """Both sides hash-partitioned by key to their owner rank with the fused shuffle, then joined
    locally (K1 + C1 + K8; reference: HashPartition x2 -> Merge -> HashJoin,
    LinqToDryad/DryadLinqQueryGen.cs:1419-1610).  The two FusedShuffle objects (receive buffers,
    pointer tables, scratch) live as long as this object and are reused by every `run`; the received
    columns are handed out as views of the receive buffers, valid until the next `run`."""

    def __init__(self, arena: SymmetricArena, ncols_left: int, ncols_right: int, cap_left: int, cap_right: int,
                 max_left: int, max_right: int):
        self.arena = arena
        self.shl = FusedShuffle(arena, [ColSpec(8, torch.int64)] * ncols_left, cap_left, max_rows_in=max_left)
        self.shr = FusedShuffle(arena, [ColSpec(8, torch.int64)] * ncols_right, cap_right, max_rows_in=max_right)

    def exchange(self, left: Sequence[torch.Tensor], right: Sequence[torch.Tensor]):
        self.shl.run(list(left), mode="hash", key_col=0)
        self.shr.run(list(right), mode="hash", key_col=0)
        ml, mr = self.shl.received_rows()[0], self.shr.received_rows()[0]
        lc = [self.shl.recv[c][0][:ml] for c in range(len(left))]
        rc = [self.shr.recv[c][0][:mr] for c in range(len(right))]
        return lc, rc

    def run(self, left: Sequence[torch.Tensor], right: Sequence[torch.Tensor]):
        """Returns (left_cols_received, right_cols_received, left_idx, right_idx): all matching pairs."""
        lc, rc = self.exchange(left, right)
        jt = K.JoinTable(rc[0])
        li, ri = jt.probe(lc[0])
        return lc, rc, li, ri

    def run_unique(self, left: Sequence[torch.Tensor], right: Sequence[torch.Tensor]):
        """Right side with unique keys: (left_cols, right_cols, right_row int32 [ml], matched uint8 [ml])."""
        lc, rc = self.exchange(left, right)
        jt = K.JoinTable(rc[0])
        ml = lc[0].numel()
        rows = torch.empty(ml, dtype=torch.int32, device=lc[0].device)
        matched = torch.empty(ml, dtype=torch.uint8, device=lc[0].device)
        N.check(N.kernels().dryad_join_probe_unique(K.ptr(jt.keys), K.ptr(jt.head), jt.cap, K.ptr(lc[0]), ml,
                                                    K.ptr(rows), K.ptr(matched), K.stream()), "join_probe_unique")
        K._count()
        return lc, rc, rows, matched


def shuffle_hash_join(arena: SymmetricArena, left: Sequence[torch.Tensor], right: Sequence[torch.Tensor],
                      cap_left: int, cap_right: int):
    """left/right: [key, payload...] int64 columns on this rank.  Both sides are hash-partitioned by
    key to their owner rank with the fused shuffle, then joined locally.  Returns
    (left_cols_received, right_cols_received, left_idx, right_idx)."""
    mark = arena.mark()
    try:
        shl = FusedShuffle(arena, [ColSpec(8, torch.int64)] * len(left), cap_left, max_rows_in=left[0].numel())
        shr = FusedShuffle(arena, [ColSpec(8, torch.int64)] * len(right), cap_right, max_rows_in=right[0].numel())
        shl.run(list(left), mode="hash", key_col=0)
        shr.run(list(right), mode="hash", key_col=0)
        ml, mr = shl.received_rows()[0], shr.received_rows()[0]
        lc = [shl.recv[c][0][:ml].clone() for c in range(len(left))]
        rc = [shr.recv[c][0][:mr].clone() for c in range(len(right))]
    finally:
        arena.release_to(mark)
    jt = K.JoinTable(rc[0].contiguous())
    li, ri = jt.probe(lc[0].contiguous())
    return lc, rc, li, ri
