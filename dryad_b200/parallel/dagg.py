"""Distributed decomposable GroupBy: local hash aggregate fused with the reduce-scatter of the
partials over NVLink (K6 + C2).

Reference plan being replaced (LinqToDryad/DryadLinqQueryGen.cs:2094-2183, 2269-2357): partial
GroupBy(seed, accumulate) → HashPartition(key) → Merge → final GroupBy(recursiveAccumulate) →
Select(finalizer), with the runtime aggregation tree of GraphManager/stagemanager/
DrDynamicAggregateManager.  On one NVSwitch domain the tree is flat: every rank folds its partial
table straight into the key owner's table with system-scope atomics on peer HBM
(`dryad_hash_agg_merge_peers`), so the shuffle, the merge and the final accumulate are ONE kernel.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..ops import core as K
from .symm import SymmetricArena


class DistributedHashAggregate:
    def __init__(self, arena: SymmetricArena, lanes: Sequence[str], local_cap: int, final_cap: int):
        self.arena, self.lanes = arena, list(lanes)
        self.world, self.rank = arena.world, arena.rank
        dev = arena.device
        stride = K.agg_lane_stride(len(self.lanes))
        fcap = max((int(final_cap) + 3) // 4 * 4, 64)
        fk, fk_off = arena.alloc((fcap,), torch.int64)
        fv, fv_off = arena.alloc((fcap, stride), torch.int64)
        self.final = K.HashAggTable(fcap, self.lanes, dev, keys=fk, vals=fv)
        self.final_cap = fcap
        self.peer_keys = [arena.peer_ptr(r, fk_off) for r in range(self.world)]
        self.peer_vals = [arena.peer_ptr(r, fv_off) for r in range(self.world)]
        self.local = K.HashAggTable(local_cap, self.lanes, dev) if self.world > 1 else None

    def run(self, keys: torch.Tensor, cols: Sequence[Optional[torch.Tensor]]):
        """keys: int64 key bits; cols: one int64/float64-bits column per lane (None for count).
        Returns (group_keys, lanes[nl][m]) for the groups THIS rank owns (hash_bucket(key, world))."""
        if self.world == 1:
            self.final.reset()
            self.final.update(keys, cols)
            return self.final.extract()
        self.final.reset()
        self.local.reset()
        self.local.update(keys, cols)
        self.arena.barrier(slot=2)                 # every owner table is reset before anyone merges
        self.local.merge_into_peers(self.world, self.peer_keys, self.peer_vals, self.final_cap)
        self.arena.barrier(slot=3)                 # all partials have landed
        if int(self.local.status.item()) == 1:
            raise K.N.NativeError("hash aggregate table full")
        return self.final.extract()
