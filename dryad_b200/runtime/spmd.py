"""SPMD job execution: one process per GPU, partition p of every stage lives on rank p.

The plan is the same physical DAG the single-process executor runs; here every rank walks it in
lock-step and executes only its own vertices.  Edges that cross partitions map to collectives
over the symmetric arena:

  CrossProduct (distributor → Merge)   fused partition+all-to-all kernel (parallel/shuffle.py) for
                                       columnar payloads; object exchange for host records
  N:1 gather (Merge to one partition)  small control-plane gather to rank 0
  1:N broadcast (side inputs)          small control-plane broadcast from rank 0
  OrderBy over byte rows               the whole sample → separators → range shuffle → sort subplan
                                       is executed by the fused `RowSorter` (parallel/dsort.py)
(The decomposable GroupBy plan runs through these edges: its HashPartition stage is the fused
shuffle; the single-kernel NVLink reduce-scatter form of the same plan is `parallel/dagg.py`,
driven directly by models/groupby.py.)

Reference roles: GraphBuilder.ConnectPointwise/ConnectCrossProduct (DryadLinqGraphManager/
GraphBuilder.cs:420-504) and the GM run loop; the reference gets these edges from files + HTTP.
"""
from __future__ import annotations

import time
from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist

from ..api.errors import DryadLinqException, ErrorCode
from ..plan.qnodes import ConnOp, DLinqQueryNode, NodeType
from .columnar import ROW_COL, ColumnBatch
from .executor import JobExecutor
from .vertex_ops import VertexEnv


def _head(n: DLinqQueryNode) -> DLinqQueryNode:
    return n.members[0] if n.type == NodeType.Super else n


def _tail(n: DLinqQueryNode) -> DLinqQueryNode:
    return n.members[-1] if n.type == NodeType.Super else n


class SpmdExecutor(JobExecutor):
    def __init__(self, context, plan, reporter=None):
        super().__init__(context, plan, reporter)
        self.rank, self.world = context.rank, context.world
        self.arena = None                      # created lazily: host-backend SPMD needs no GPU arena
        self._skip: set = set()
        self._fused: Dict[int, Any] = {}

    # ---- pattern detection ---------------------------------------------------------------------
    def _find_row_sort_patterns(self):
        """OrderBy(RowKey) fed by Merge ← RangePartition ← (X, RangeSampler ← Merge ← Sampler ← X):
        the OrderBy stage is executed by RowSorter directly on X's local partition."""
        from ..api.rows import RowKey
        for n in self.plan.nodes:
            members = n.members if n.type == NodeType.Super else [n]
            ob = next((m for m in members if m.type == NodeType.OrderBy), None)
            if ob is None or not isinstance(getattr(ob.params["key"], "fn", None), RowKey):
                continue
            if _head(n).type != NodeType.Merge or not n.children or n.children[0][1] != ConnOp.CrossProduct:
                continue
            rp = n.children[0][0]
            if _tail(rp).type != NodeType.RangePartition or len(rp.children) != 2:
                continue
            base = rp.children[0][0]
            chain = []
            cur = rp.children[1][0]
            while cur is not base and cur.children:
                chain.append(cur)
                cur = cur.children[0][0]
            if cur is not base:
                continue
            self._fused[n.uid] = ("rowsort", base, ob, members[members.index(ob) + 1:])
            self._skip.update(x.uid for x in chain)
            self._skip.add(rp.uid)

    # ---- run -----------------------------------------------------------------------------------
    def run(self):
        ctx = self.context
        if any(n.partition_count not in (1, self.world) for n in self.plan.nodes):
            raise DryadLinqException(ErrorCode.Internal,
                                     "SPMD execution needs every stage to have 1 or world partitions")
        self._find_row_sort_patterns()
        # GPU backend: distributor stages (HashPartition / RangePartition into `world` ports) run
        # the fused partition + all-to-all kernel over the symmetric arena instead of emitting ports
        if getattr(self.backend, "name", "") == "gpu":
            self.backend.spmd_exchange = True
        rep = self.reporter
        rep.job_start(self.plan)
        t_job = time.perf_counter()
        for node in self.plan.nodes:
            self.out[node.uid] = [None] * node.partition_count
            if node.uid in self._skip:
                continue
            mine = self.rank if node.partition_count == self.world else (0 if self.rank == 0 else None)
            t0 = time.perf_counter()
            import contextlib
            span = self.timeline.span(f"{node.name}#{node.uid}[{mine}]", stage=node.uid, part=mine, rank=self.rank) \
                if (self.timeline is not None and mine is not None) else contextlib.nullcontext()
            if node.uid in self._fused:
                with span:
                    ports = self._run_fused(node)
            elif mine is None:
                self._participate_only(node)
                continue
            else:
                ins = self._spmd_inputs(node, mine)
                env = VertexEnv(ctx, node.uid, mine, 0)
                with span:
                    ports = self.backend.run_vertex(node, mine, ins, env)
            if mine is not None:
                self.out[node.uid][mine] = ports
                rep.vertex_event(node, mine, 0, "completed", seconds=time.perf_counter() - t0)
        results = self._commit_local()
        self._flush_timeline()
        rep.job_stop(True, "", seconds=time.perf_counter() - t_job)
        return results

    def _participate_only(self, node):
        """Ranks without a vertex in a single-partition stage still take part in its gathers."""
        for child, op in node.children:
            if child.partition_count == self.world:
                self._gather_to_zero(child)

    # ---- edges -----------------------------------------------------------------------------------
    def _gather_to_zero(self, child) -> Optional[List]:
        local = self.out[child.uid][self.rank]
        payload = self.backend.to_records(local[0]) if local is not None else []
        bucket: Optional[List] = [None] * self.world if self.rank == 0 else None
        dist.gather_object(payload, bucket, dst=0)
        return bucket

    def _spmd_inputs(self, node, i) -> List[List]:
        ins: List[List] = []
        head = _head(node)
        if head.type == NodeType.SlidingWindow and node.partition_count == self.world:
            # C6 neighbour shift: a partition only needs the w-1 records that precede it, so every
            # rank publishes just its tail (a few records) instead of exchanging partitions
            # (reference: HashPartition(x => x.Index) of the boundary records,
            # LinqToDryad/DryadLinqQueryGen.cs:3273-3280)
            w = head.params["window"]
            child = node.children[0][0]
            own = self.out[child.uid][self.rank][0]
            recs = self.backend.to_records(own)
            tails: List = [None] * self.world
            dist.all_gather_object(tails, list(recs[-(w - 1):]) if w > 1 else [])
            halo: List = []
            s = self.rank - 1
            while s >= 0 and len(halo) < w - 1:
                seg = tails[s]
                take = (w - 1) - len(halo)
                halo = list(seg[-take:]) + halo if take > 0 and seg else halo
                s -= 1
            return [[own], [halo if self.rank > 0 else []]]
        for child, op in node.children:
            cpc = child.partition_count
            if op == ConnOp.CrossProduct:
                ports = self.out[child.uid][self.rank] if cpc == self.world else (self.out[child.uid][0] if self.rank == 0 else None)
                if ports is not None and len(ports) == 1 and getattr(ports[0], "_exchanged", False):
                    ins.append([ports[0]])                # the distributor already ran the fused all-to-all
                    continue
                recs = [self.backend.to_records(p) for p in ports] if ports is not None else None
                everyone: List = [None] * self.world
                dist.all_gather_object(everyone, recs)
                ins.append([everyone[s][i] for s in range(self.world) if everyone[s] is not None])
            elif cpc == node.partition_count:
                ins.append([self.out[child.uid][i][0]])
            elif cpc == 1:                                    # 1:N broadcast from rank 0
                box = [self.backend.to_records(self.out[child.uid][0][0]) if self.rank == 0 else None]
                dist.broadcast_object_list(box, src=0)
                ins.append([box[0]])
            elif node.partition_count == 1:                   # N:1 gather (only rank 0 gets here)
                ins.append(self._gather_to_zero(child))
            else:
                raise DryadLinqException(ErrorCode.Internal, "unsupported SPMD edge")
        return ins

    # ---- fused stages ----------------------------------------------------------------------------
    def _run_fused(self, node):
        kind, base, ob, rest = self._fused[node.uid]
        from ..parallel.dsort import RowSorter
        batch: ColumnBatch = self.out[base.uid][self.rank][0]
        rows = batch.cols[ROW_COL]
        if rows.device.type != "cuda":
            rows = rows.to(self.backend.device, non_blocking=True)
        rk = ob.params["key"].fn
        nmax = torch.tensor([rows.shape[0]], dtype=torch.int64, device=rows.device)
        dist.all_reduce(nmax, op=dist.ReduceOp.MAX)           # control plane: size the receive side
        sorter = self.context._row_sorter(int(nmax.item()), rows.shape[1], rk.offset, max(8, rk.length),
                                          bool(ob.params["descending"]))
        out, m = sorter.sort(rows.contiguous())
        res = ColumnBatch.rows(out[:m], rk.offset, rk.length)
        ports = [res]
        env = VertexEnv(self.context, node.uid, self.rank, 0)
        for mnode in rest:                                     # e.g. the OutputTable member
            ports = self.backend._run_one(mnode, self.rank, [[ports[0]]], env)
        return ports

    def _commit_local(self):
        results = {}
        for root in self.plan.roots:
            node = _tail(root)
            mine = self.rank if root.partition_count == self.world else 0
            payload = self.out[root.uid][mine]
            parts = [payload[0]] if payload is not None else [[]]
            table = self.context._commit_table(None, parts, root.schema, root.info, None, True)
            results[node.params["qexpr_id"]] = table
        return results
