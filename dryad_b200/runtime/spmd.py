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
  decomposable GroupBy                 partial GroupBy → HashPartition → Merge → final GroupBy runs as
                                       ONE DistributedHashAggregate (local table + NVLink reduce-scatter
                                       into the key owners' tables, parallel/dagg.py)
  Join(broadcast=True)                 to-one Merge → Tee → CrossProduct → Merge feeding the Join runs as
                                       BroadcastJoin (NVLS / P2P replication fused with the build) and the
                                       generated probe kernel (parallel/djoin.py)
Columnar payloads never travel as pickled objects: the remaining 1:N / N:1 / all-to-all edges move
tensors (padded all_gather / broadcast).  `object_collectives` counts the object exchanges a job
needed (host records only).

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
from .spmd_control import StageAborted
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
        self._cp = None
        self.object_collectives = 0            # data-carrying *_object collectives (host records)
        self.tensor_collectives = 0            # columnar edges moved as tensors

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

    def _find_groupby_patterns(self):
        """[... | GroupBy(partial_acc) | HashPartition] ×→ [Merge | GroupBy(final_acc) | ...] with GPU
        lanes: producer and consumer are executed as one DistributedHashAggregate
        (reference plan: LinqToDryad/DryadLinqQueryGen.cs:2094-2183)."""
        for c in self.plan.nodes:
            cm = c.members if c.type == NodeType.Super else [c]
            if len(cm) < 2 or cm[0].type != NodeType.Merge or cm[1].type != NodeType.GroupBy:
                continue
            if cm[1].params.get("mode") != "final_acc" or len(c.children) != 1 or c.children[0][1] != ConnOp.CrossProduct:
                continue
            n = c.children[0][0]
            nm = n.members if n.type == NodeType.Super else [n]
            if len(nm) < 2 or nm[-1].type != NodeType.HashPartition or nm[-2].type != NodeType.GroupBy:
                continue
            gb = nm[-2]
            dec = gb.params.get("decomp")
            if gb.params.get("mode") != "partial_acc" or dec is None or gb.params.get("elem") is not None:
                continue
            if dec.gpu_lanes() is None or gb.params.get("comparer") is not None:
                continue
            if n.partition_count != self.world or c.partition_count != self.world:
                continue
            if sum(1 for x in self.plan.nodes for ch, _ in x.children if ch is n) != 1:
                continue                                   # the partials feed something else too
            self._fused[n.uid] = ("gb_producer", nm[:-2], gb, c)
            self._fused[c.uid] = ("gb_consumer", n, cm[1], cm[2:])

    def _find_broadcast_join_patterns(self):
        """Join(broadcast=True) whose inner side is Merge(world) ×← Tee ← Merge(1) ← X: X's local
        partition goes straight into BroadcastJoin (LinqToDryad/DryadLinqQueryGen.cs:2957-3001)."""
        for j in self.plan.nodes:
            jm = j.members if j.type == NodeType.Super else [j]
            jn = jm[0]
            if jn.type != NodeType.Join or not jn.params.get("broadcast") or len(j.children) != 2:
                continue
            bc = j.children[1][0]
            if bc.type != NodeType.Merge or not bc.children or bc.children[0][1] != ConnOp.CrossProduct:
                continue
            tee = bc.children[0][0]
            if _tail(tee).type != NodeType.Tee or tee.partition_count != 1 or not tee.children:
                continue
            chain = [bc, tee]
            one = tee if tee.type == NodeType.Super else None
            src = tee.children[0][0]
            if one is None:
                if src.type != NodeType.Merge or src.partition_count != 1 or not src.children:
                    continue
                chain.append(src)
                src = src.children[0][0]
            elif _head(tee).type != NodeType.Merge:
                continue
            if src.partition_count != self.world:
                continue
            if any(sum(1 for x in self.plan.nodes for ch, _ in x.children if ch is y) != 1 for y in chain):
                continue
            self._fused[j.uid] = ("bjoin", src, jn, jm[1:])
            self._skip.update(x.uid for x in chain)

    # ---- run -----------------------------------------------------------------------------------
    def _control(self):
        ctx = self.context
        cp = ctx.__dict__.get("_control_plane")
        if cp is None:
            from .spmd_control import ControlPlane
            dev = self.backend.device if getattr(self.backend, "name", "") == "gpu" else torch.device("cpu")
            cp = ControlPlane(ctx, dev)
            ctx.__dict__["_control_plane"] = cp
        return cp

    def _sync(self) -> None:
        """Agreement point, called before every collective of a stage attempt (see spmd_control)."""
        if self._cp is not None and not self._cp.agree(True):
            raise StageAborted()

    def _build_gm(self):
        """One GM vertex per executed plan stage (an SPMD stage is a gang), chained in plan order on
        top of the plan's own edges so the manager hands the stages out in an order every rank can
        follow with the inputs it retained."""
        from ..sched.gm import GraphManager
        ctx = self.context
        gm = GraphManager(max_failures=ctx.max_failures_per_vertex, speculative=False)
        self._gm_node = {}
        prev = None
        sid = {}
        for node in self.plan.nodes:
            if node.uid in self._skip:
                continue
            sid[node.uid] = gm.add_stage(f"{node.name}#{node.uid}", 1)
            self._gm_node[sid[node.uid]] = node
            if prev is not None:
                gm.connect(prev, sid[node.uid])
            prev = sid[node.uid]
        return gm

    def run(self):
        ctx = self.context
        if any(n.partition_count not in (1, self.world) for n in self.plan.nodes):
            raise DryadLinqException(ErrorCode.Internal,
                                     "SPMD execution needs every stage to have 1 or world partitions")
        self._find_row_sort_patterns()
        if getattr(self.backend, "name", "") == "gpu" and self.world > 1:
            self._find_groupby_patterns()
            self._find_broadcast_join_patterns()
        # GPU backend: distributor stages (HashPartition / RangePartition into `world` ports) run
        # the fused partition + all-to-all kernel over the symmetric arena instead of emitting ports
        if getattr(self.backend, "name", "") == "gpu":
            self.backend.spmd_exchange = True
        rep = self.reporter
        rep.job_start(self.plan)
        t_job = time.perf_counter()
        for node in self.plan.nodes:
            self.out[node.uid] = [None] * node.partition_count
        self._cp = self._control() if self.world > 1 else None
        ctx.__dict__["_spmd_sync"] = self._sync
        nodes = {n.uid: n for n in self.plan.nodes}
        try:
            if self._cp is None:
                for node in self.plan.nodes:
                    if node.uid not in self._skip:
                        self._run_stage(node, 0)
                ok, message = True, ""
            else:
                ok, message = self._run_controlled(nodes, rep)
        finally:
            ctx.__dict__.pop("_spmd_sync", None)
        ctx.__dict__["last_object_collectives"] = self.object_collectives
        ctx.__dict__["last_tensor_collectives"] = self.tensor_collectives
        if not ok:
            self.out.clear()                               # nothing is committed
            rep.job_stop(False, message)
            raise DryadLinqException(ErrorCode.DidNotCompleteSuccessfully, message)
        results = self._commit_local()
        self._flush_timeline()
        rep.job_stop(True, "", seconds=time.perf_counter() - t_job)
        return results

    def _run_controlled(self, nodes, rep):
        """The job under the control plane: rank 0 asks the native graph manager for the next stage
        attempt and posts its launch descriptor into every rank's mailbox; every rank (rank 0
        included) runs the attempt and reports; a failed attempt is re-run by all ranks from the
        inputs they retained; when the failure budget of a stage is spent (or the error is
        deterministic) every rank receives `abort`; outputs are committed only on `commit`."""
        import traceback
        from .command import CMD_START, VertexCommand
        from .spmd_control import terminate_command
        ctx, cp = self.context, self._cp
        gm = self._build_gm() if self.rank == 0 else None
        self.gm = gm
        if gm is not None:
            gm.start()
        t_job = time.perf_counter()
        self.stage_attempts = 0
        while True:
            if gm is not None:                             # ---- graph manager (rank 0) ----
                nxt = gm.next_ready(time.perf_counter() - t_job)
                if nxt is None:
                    done = gm.state == "Succeeded"
                    cp.post_command(terminate_command(done, "" if done else f"job ended in state {gm.state}: {gm.error()}"))
                else:
                    v, version = nxt
                    node = self._gm_node[gm.vertex(v)["stage"]]
                    cp.post_command(VertexCommand(CMD_START, node.uid, version, args=[node.name, str(self.world)]))
            cmd = cp.next_command()                        # ---- worker (every rank) ----
            if cmd.command != CMD_START:
                commit = bool(cmd.args) and cmd.args[0] == "commit"
                return commit, (cmd.args[1] if len(cmd.args) > 1 else "")
            node, version = nodes[cmd.vertex_id], cmd.version
            self.stage_attempts += 1
            t0 = time.perf_counter()
            ok, text, deterministic = True, "", False
            try:
                if ctx.fault_injector is not None:
                    ctx.fault_injector(node, self.rank, version)
                self._run_stage(node, version)
                if not cp.agree(True):                     # end-of-attempt agreement
                    ok, text = False, "aborted: another rank failed"
            except StageAborted:
                ok, text = False, "aborted: another rank failed"
            except Exception as ex:  # noqa: BLE001  (vertex code is user code)
                cp.agree(False)                            # pairs with the healthy ranks' next agreement point
                ok = False
                text = "".join(traceback.format_exception_only(type(ex), ex)).strip()
                deterministic = isinstance(ex, DryadLinqException) and not getattr(ex, "transient", False)
                if deterministic and ctx.fail_fast_on_deterministic_errors:
                    text = "deterministic: " + text
            if not ok:
                self.out[node.uid] = [None] * node.partition_count
            rep.vertex_event(node, self.rank, version, "completed" if ok else "failed",
                             **({"seconds": time.perf_counter() - t0} if ok else {"error": text}))
            cp.report(node.uid, version, ok, text)
            if gm is not None:                             # ---- graph manager: account the attempt ----
                all_ok, err = cp.collect(node.uid, version)
                v = next(k for k, n_ in self._gm_node.items() if n_ is node)
                gv = gm.stage_vertices(v)[0]
                if all_ok:
                    gm.completed(gv, version, time.perf_counter() - t0)
                else:
                    rc = gm.failed(gv, version, err or "stage attempt failed")
                    if rc == 1 or "deterministic: " in err or "rank lost" in err:
                        msg = gm.error() or err
                        if "deterministic: " in err:
                            msg = err.replace("deterministic: ", "")
                        cp.post_command(terminate_command(False, msg))
                        gm = None                          # the next command every rank sees is `abort`

    def _run_stage(self, node, version: int):
        """One attempt of one plan stage on this rank."""
        ctx, rep = self.context, self.reporter
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
            return
        else:
            ins = self._spmd_inputs(node, mine)
            env = VertexEnv(ctx, node.uid, mine, version)
            with span:
                ports = self.backend.run_vertex(node, mine, ins, env)
        if mine is not None:
            self.out[node.uid][mine] = ports
            if self._cp is None:
                rep.vertex_event(node, mine, version, "completed", seconds=time.perf_counter() - t0)

    def _participate_only(self, node):
        """Ranks without a vertex in a single-partition stage still take part in its collectives: this
        mirrors, edge by edge, what rank 0 does in `_spmd_inputs` (same agreement points, same
        collectives in the same order), so no rank waits in a collective the others never enter."""
        for child, op in node.children:
            cpc = child.partition_count
            if op == ConnOp.CrossProduct:
                ports = self.out[child.uid][self.rank] if cpc == self.world else None
                if ports is not None and len(ports) == 1 and getattr(ports[0], "_exchanged", False):
                    continue
                if cpc == 1 and _tail(child).type == NodeType.Tee and _tail(child).params.get("replicate") \
                        and self._all_columnar([], child.schema):
                    self._allgather_batches(None, child.schema, False)
                    continue
                recs = [self.backend.to_records(p) for p in ports] if ports is not None else None
                everyone: List = [None] * self.world
                self._sync()
                dist.all_gather_object(everyone, recs)
                self.object_collectives += 1
            elif cpc == node.partition_count:
                continue
            elif cpc == self.world:
                self._gather_to_zero(child)

    # ---- edges -----------------------------------------------------------------------------------
    _DT = {"i": torch.int64, "f": torch.float64, "b": torch.bool}

    def _all_columnar(self, payloads, schema) -> bool:
        """Collective decision (one tiny all-reduce): every rank that holds a payload for this edge
        holds it as a ColumnBatch with a known numeric schema.  A rank whose vertex fell back to host
        records makes the whole edge take the object path, so no rank waits in the wrong collective."""
        from .gpu_backend import _numeric_schema
        self._sync()
        ok = _numeric_schema(schema) and all(isinstance(p, ColumnBatch) and not p.is_rows and p.acc_lanes is None
                                             for p in payloads if p is not None)
        dev = self.backend.device if getattr(self.backend, "name", "") == "gpu" else torch.device("cpu")
        flag = torch.tensor([1 if ok else 0], dtype=torch.int64, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        return bool(int(flag.item()))

    def _allgather_batches(self, mine: Optional[ColumnBatch], schema, tuple_like=False) -> List[ColumnBatch]:
        """All-gather of one ColumnBatch per rank as tensors: sizes first, then every column padded
        to the largest partition (one all_gather_into_tensor per column)."""
        names = list(schema)
        dev = self.backend.device if getattr(self.backend, "name", "") == "gpu" else torch.device("cpu")
        n = mine.n if mine is not None else 0
        sizes = torch.zeros(self.world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(sizes, torch.tensor([n], dtype=torch.int64, device=dev))
        sz = sizes.tolist()
        cap = max(max(sz), 1)
        cols: Dict[str, torch.Tensor] = {}
        for k in names:
            dt = self._DT[schema[k]]
            wire = torch.uint8 if dt == torch.bool else dt
            buf = torch.zeros(cap, dtype=wire, device=dev)
            if n:
                buf[:n] = mine.cols[k].to(dev).to(wire)
            out = torch.empty(self.world * cap, dtype=wire, device=dev)
            dist.all_gather_into_tensor(out, buf)
            cols[k] = out
        self.tensor_collectives += 1
        res = []
        for r in range(self.world):
            d = {k: (cols[k][r * cap:r * cap + sz[r]].to(torch.bool) if self._DT[schema[k]] == torch.bool
                     else cols[k][r * cap:r * cap + sz[r]]) for k in names}
            res.append(ColumnBatch(d, sz[r], tuple_like))
        return res

    def _halo_p2p(self, own: ColumnBatch, w: int, schema) -> Optional[ColumnBatch]:
        """C6 neighbour shift over NVLink: every rank stores the last w-1 rows of its partition into
        the NEXT rank's halo slot in the symmetric arena (plain P2P stores), one device barrier, and
        the receiver reads its slot locally.  Applies when every partition holds at least w-1 rows
        (otherwise a halo spans several ranks and the all-gather form is used).  GPU backend only."""
        if getattr(self.backend, "name", "") != "gpu" or own.device.type != "cuda":
            return None
        dev = self.backend.device
        nmin = torch.tensor([own.n], dtype=torch.int64, device=dev)
        dist.all_reduce(nmin, op=dist.ReduceOp.MIN)
        if int(nmin.item()) < w - 1:
            return None
        from ..parallel.symm import tensor_from_ptr
        ctx = self.context
        arena = ctx._spmd_arena()
        names = list(schema)
        cache = ctx.__dict__.setdefault("_halo_cache", {})
        sig = (len(names), w)
        slot = cache.get(sig)
        if slot is None:
            buf, off = arena.alloc((len(names), w - 1), torch.int64)
            slot = cache[sig] = (buf, off)
        buf, off = slot
        h = w - 1
        arena.barrier(slot=2)                          # the neighbour has consumed its previous halo
        if self.rank + 1 < self.world:
            for c, k in enumerate(names):
                v = own.cols[k][own.n - h:]
                wire = v.view(torch.int64) if v.dtype == torch.float64 else v.to(torch.int64)
                dst = tensor_from_ptr(arena.peer_ptr(self.rank + 1, off + c * h * 8), h * 8, dev).view(torch.int64)
                dst.copy_(wire, non_blocking=True)
        arena.barrier(slot=3)                          # every halo has landed
        self.p2p_halos = getattr(self, "p2p_halos", 0) + 1
        ctx.__dict__["last_p2p_halos"] = self.p2p_halos
        if self.rank == 0:
            return own.slice(0, 0)
        cols = {}
        for c, k in enumerate(names):
            raw = buf[c].clone()
            kind = schema[k]
            cols[k] = raw.view(torch.float64) if kind == "f" else (raw != 0 if kind == "b" else raw)
        return ColumnBatch(cols, h, getattr(own, "tuple_like", False))

    def _pipe_to_zero(self, payload: Optional[ColumnBatch], schema) -> Optional[List]:
        """N:1 merge of columnar partitions as STREAMED rank-to-rank channels (C7): every rank r > 0
        streams its partition to rank 0 in blocks — `channel/pipe.py` (point-to-point sends on a side
        stream, bounded depth) or, with context.spmd_pipe == "ring", the HBM ring FIFO of
        `parallel/ring.py` (P2P stores + release/acquire flags, no NCCL) — while rank 0 receives them in
        rank order.  Only rank 0 ends up with the data (the all-gather form gave it to everyone)."""
        names = sorted(schema)
        dts = {k: self._DT[schema[k]] for k in names}
        dev = self.backend.device if getattr(self.backend, "name", "") == "gpu" else torch.device("cpu")
        tl = getattr(payload, "tuple_like", False) if payload is not None else False
        mode = getattr(self.context, "spmd_pipe", "nccl")
        self.tensor_collectives += 1
        self.streamed_edges = getattr(self, "streamed_edges", 0) + 1
        self.context.__dict__["last_streamed_edges"] = self.streamed_edges
        if mode == "ring" and dev.type == "cuda":
            from ..parallel.ring import HbmRing, ring_recv_batch, ring_send_batch
            ctx = self.context
            rings = ctx.__dict__.get("_gather_rings")
            if rings is None:
                arena = ctx._spmd_arena()
                rings = ctx.__dict__["_gather_rings"] = [None] + [HbmRing(arena, r, 0, slots=4, slot_bytes=1 << 20)
                                                                  for r in range(1, self.world)]
            if self.rank == 0:
                parts = [payload if payload is not None else ColumnBatch({k: torch.empty(0, dtype=dts[k], device=dev) for k in names}, 0)]
                for r in range(1, self.world):
                    cols = ring_recv_batch(rings[r], names, dts)
                    parts.append(ColumnBatch({k: cols[k] for k in schema}, None, tl))
                return parts
            ring_send_batch(rings[self.rank], payload, names)
            rings[self.rank].check()
            return None
        from ..channel.pipe import PipeReader, PipeWriter, recv_batch, send_batch
        if self.rank == 0:
            parts = [payload if payload is not None else ColumnBatch({k: torch.empty(0, dtype=dts[k], device=dev) for k in names}, 0)]
            for r in range(1, self.world):
                cols = recv_batch(PipeReader(r, dev, tag=7), names)
                parts.append(ColumnBatch({k: cols[k].to(dts[k]) if cols[k].numel() == 0 else cols[k] for k in schema}, None, tl))
            return parts
        if payload is None:
            payload = ColumnBatch({k: torch.empty(0, dtype=dts[k], device=dev) for k in names}, 0)
        send_batch(PipeWriter(0, dev, tag=7), payload)
        return None

    def _gather_to_zero(self, child) -> Optional[List]:
        local = self.out[child.uid][self.rank]
        payload = local[0] if local is not None else None
        if self._all_columnar([payload], child.schema):
            return self._pipe_to_zero(payload, child.schema)
        recs = self.backend.to_records(payload) if payload is not None else []
        bucket: Optional[List] = [None] * self.world if self.rank == 0 else None
        self._sync()
        dist.gather_object(recs, bucket, dst=0)
        self.object_collectives += 1
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
            if w > 1 and self._all_columnar([own], child.schema):
                halo_p2p = self._halo_p2p(own, w, child.schema)
                if halo_p2p is not None:
                    return [[own], [halo_p2p]]
                tail = own.slice(max(0, own.n - (w - 1)), own.n)
                tails = self._allgather_batches(tail, child.schema, getattr(own, "tuple_like", False))
                halo_parts: List[ColumnBatch] = []
                have = 0
                s = self.rank - 1
                while s >= 0 and have < w - 1:
                    seg = tails[s]
                    take = min(seg.n, (w - 1) - have)
                    if take:
                        halo_parts.insert(0, seg.slice(seg.n - take, seg.n))
                        have += take
                    s -= 1
                halo = ColumnBatch.concat(halo_parts) if halo_parts else own.slice(0, 0)
                return [[own], [halo if self.rank > 0 else own.slice(0, 0)]]
            recs = self.backend.to_records(own)
            tails_o: List = [None] * self.world
            self._sync()
            dist.all_gather_object(tails_o, list(recs[-(w - 1):]) if w > 1 else [])
            self.object_collectives += 1
            halo_r: List = []
            s = self.rank - 1
            while s >= 0 and len(halo_r) < w - 1:
                seg = tails_o[s]
                take = (w - 1) - len(halo_r)
                halo_r = list(seg[-take:]) + halo_r if take > 0 and seg else halo_r
                s -= 1
            return [[own], [halo_r if self.rank > 0 else []]]
        if head.type == NodeType.ForkBranch:
            # branch k of a Fork reads output port k of the Fork vertex of the same partition
            child = node.children[0][0]
            if child.partition_count == node.partition_count:
                return [[self.out[child.uid][i][head.params["index"]]]]
        if head.type == NodeType.Reverse and node.partition_count == self.world and node.children[0][0].partition_count == self.world:
            # partition i of the result is partition world-1-i of the input, reversed locally by the vertex
            # (DryadLinqQueryGen.cs:2731-2810): a permutation of whole partitions between ranks
            parts = self._gather_all(node.children[0][0])
            return [[parts[self.world - 1 - self.rank]]] + [[self.out[c.uid][i][0]] for c, _ in node.children[1:]]
        if head.type == NodeType.Concat and node.partition_count == self.world:
            # partition i of the result is one whole partition of one child, by position
            acc = 0
            picked = None
            for child, _ in node.children:
                parts = self._gather_all(child)
                if picked is None and i < acc + child.partition_count:
                    picked = parts[i - acc]
                acc += child.partition_count
            return [[picked if picked is not None else []]]
        if head.type == NodeType.Zip and node.partition_count == self.world and len(node.children) == 2 \
                and node.children[0][0].partition_count == self.world:
            got = self._zip_inputs(node)
            if got is not None:
                return got
        for child, op in node.children:
            cpc = child.partition_count
            if op == ConnOp.CrossProduct:
                ports = self.out[child.uid][self.rank] if cpc == self.world else (self.out[child.uid][0] if self.rank == 0 else None)
                if ports is not None and len(ports) == 1 and getattr(ports[0], "_exchanged", False):
                    ins.append([ports[0]])                # the distributor already ran the fused all-to-all
                    continue
                if cpc == 1 and _tail(child).type == NodeType.Tee and _tail(child).params.get("replicate") \
                        and self._all_columnar(ports or [], child.schema):
                    # replicated Tee ports are the same batch: broadcast it once as tensors
                    got = self._allgather_batches(ports[0] if ports else None, child.schema,
                                                  getattr(ports[0], "tuple_like", False) if ports else False)
                    ins.append([got[0]])
                    continue
                recs = [self.backend.to_records(p) for p in ports] if ports is not None else None
                everyone: List = [None] * self.world
                self._sync()
                dist.all_gather_object(everyone, recs)
                self.object_collectives += 1
                ins.append([everyone[s][i] for s in range(self.world) if everyone[s] is not None])
            elif cpc == node.partition_count:
                ins.append([self.out[child.uid][i][0]])
            elif cpc == 1:                                    # 1:N broadcast from rank 0
                payload = self.out[child.uid][0][0] if self.rank == 0 else None
                if self._all_columnar([payload], child.schema):
                    got = self._allgather_batches(payload, child.schema, getattr(payload, "tuple_like", False))
                    ins.append([got[0]])
                    continue
                box = [self.backend.to_records(payload) if self.rank == 0 else None]
                self._sync()
                dist.broadcast_object_list(box, src=0)
                self.object_collectives += 1
                ins.append([box[0]])
            elif node.partition_count == 1:                   # N:1 gather (only rank 0 gets here)
                ins.append(self._gather_to_zero(child))
            else:
                raise DryadLinqException(ErrorCode.Internal, "unsupported SPMD edge")
        return ins

    def _gather_all(self, child) -> List:
        """Every partition of `child` on every rank (collective): tensors when the stage is columnar on all
        ranks, records otherwise.  For the operators that move WHOLE partitions by position (Reverse, Concat)."""
        cpc = child.partition_count
        own = self.out[child.uid][self.rank][0] if cpc == self.world else (self.out[child.uid][0][0] if self.rank == 0 else None)
        if self._all_columnar([own], child.schema):
            parts = self._allgather_batches(own, child.schema, getattr(own, "tuple_like", False) if own is not None else False)
            return parts if cpc == self.world else [parts[0]]
        recs = list(self.backend.to_records(own)) if own is not None else None
        box: List = [None] * self.world
        self._sync()
        dist.all_gather_object(box, recs)
        self.object_collectives += 1
        return [b if b is not None else [] for b in box] if cpc == self.world else [box[0] or []]

    def _zip_inputs(self, node) -> Optional[List[List]]:
        """Zip pairs the i-th record of the left side with the i-th record of the right side GLOBALLY, so
        the right side is re-aligned to the left side's partition boundaries (DryadLinqQueryGen.cs:2626-2717).
        Every rank learns every rank's (left, right) counts (one tiny all-gather).  When the boundaries
        already coincide nothing moves; otherwise the right side is gathered — as tensors when it is
        columnar on every rank, as records otherwise — and each rank keeps the slice under its left range."""
        lchild, rchild = node.children[0][0], node.children[1][0]
        left = self.out[lchild.uid][self.rank][0]
        rpc = rchild.partition_count
        right = self.out[rchild.uid][self.rank][0] if rpc == self.world else (self.out[rchild.uid][0][0] if self.rank == 0 else None)
        dev = self.backend.device if getattr(self.backend, "name", "") == "gpu" else torch.device("cpu")
        mine = torch.tensor([self.backend.count(left), self.backend.count(right) if right is not None else 0],
                            dtype=torch.int64, device=dev)
        everyone = [torch.zeros(2, dtype=torch.int64, device=dev) for _ in range(self.world)]
        self._sync()
        dist.all_gather(everyone, mine)
        counts = torch.stack(everyone).cpu().tolist()
        if rpc == self.world and all(a == b for a, b in counts):
            return [[left], [right]]                         # same boundaries on both sides: purely local
        start = sum(c[0] for c in counts[:self.rank])
        need = counts[self.rank][0]
        if self._all_columnar([right], rchild.schema):
            parts = self._allgather_batches(right, rchild.schema, getattr(right, "tuple_like", False) if right is not None else False)
            whole = ColumnBatch.concat(parts)
            return [[left], [whole.slice(min(start, whole.n), min(start + need, whole.n))]]
        recs = list(self.backend.to_records(right)) if right is not None else []
        box: List = [None] * self.world
        self._sync()
        dist.all_gather_object(box, recs)
        self.object_collectives += 1
        flat = [x for b in box for x in (b or [])]
        return [[self.backend.to_records(left)], [flat[start:start + need]]]

    # ---- fused stages ----------------------------------------------------------------------------
    def _run_fused(self, node):
        kind = self._fused[node.uid][0]
        if kind == "rowsort":
            return self._run_rowsort(node)
        if kind == "gb_producer":
            return self._run_gb_producer(node)
        if kind in ("gb_consumer", "plain"):
            return self._run_gb_consumer(node)
        return self._run_bjoin(node)

    def _run_rowsort(self, node):
        kind, base, ob, rest = self._fused[node.uid]
        batch: ColumnBatch = self.out[base.uid][self.rank][0]
        rows = batch.cols[ROW_COL]
        if rows.device.type != "cuda":
            rows = rows.to(self.backend.device, non_blocking=True)
        rk = ob.params["key"].fn
        self._sync()
        nmax = torch.tensor([rows.shape[0]], dtype=torch.int64, device=rows.device)
        dist.all_reduce(nmax, op=dist.ReduceOp.MAX)           # control plane: size the receive side
        sorter = self.context._row_sorter(int(nmax.item()), rows.shape[1], rk.offset, max(8, rk.length),
                                          bool(ob.params["descending"]))
        out, m = sorter.sort(rows.contiguous())
        res = ColumnBatch.rows(out[:m], rk.offset, rk.length)
        return self._run_rest([res], node, rest)

    def _run_rest(self, ports, node, rest):
        env = VertexEnv(self.context, node.uid, self.rank, 0)
        for mnode in rest:                                     # e.g. Select_FinalReduce, OutputTable
            ports = self.backend._run_one(mnode, self.rank, [[ports[0]]], env)
        return ports

    # -- decomposable GroupBy as one DistributedHashAggregate ------------------------------------
    def _run_gb_producer(self, node):
        """Runs the members ahead of the partial GroupBy, then only PREPARES the aggregate inputs
        (key bits + one column per lane).  All ranks agree (one all-reduce) whether the fused form
        applies; if not, the GroupBy and HashPartition members run as planned."""
        from . import gpu_backend as G
        kind, prefix, gb, consumer = self._fused[node.uid]
        be = self.backend
        env = VertexEnv(self.context, node.uid, self.rank, 0)
        cur = self._spmd_inputs(node, self.rank)
        for mnode in prefix:
            out = be._run_one(mnode, self.rank, cur, env)
            cur = [[out[0]]]
        pre = NotImplemented
        if all(isinstance(p, ColumnBatch) and not p.is_rows for ps in cur for p in ps):
            try:
                pre = G._group_by_inputs(be, gb, be._cat(cur[0]))
            except G.CompileError:
                pre = NotImplemented
        ok = pre is not NotImplemented and not G._has_reserved_key(pre[0])
        self._sync()
        flag = torch.tensor([1 if ok else 0], dtype=torch.int64, device=be.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()):
            self._gb_ready = (node.uid, pre)
            return [None]
        # planned path: partial table, fused hash shuffle, final table
        self._fused[consumer.uid] = ("plain",)
        members = (node.members if node.type == NodeType.Super else [node])[len(prefix):]
        for mnode in members:
            out = be._run_one(mnode, self.rank, cur, env)
            cur = [[out[0]]]
        return out

    def _dagg(self, lanes):
        """One DistributedHashAggregate per lane signature for the life of the context (tables in
        the symmetric arena, peer pointer tables built once)."""
        from ..ops import core as K
        from ..parallel.dagg import DistributedHashAggregate
        ctx = self.context
        cache = ctx.__dict__.setdefault("_dagg_cache", {})
        hint = int(ctx.group_capacity_hint)
        sig = (tuple(lanes), hint)
        agg = cache.get(sig)
        if agg is None:
            fcap = K.agg_capacity(int(1.25 * hint) // self.world + 1024)
            agg = DistributedHashAggregate(ctx._spmd_arena(), list(lanes), K.agg_capacity(hint), fcap)
            cache[sig] = agg
        return agg

    def _run_gb_consumer(self, node):
        from . import gpu_backend as G
        info = self._fused[node.uid]
        if info[0] == "plain":                                 # the producer took the planned path
            ins = self._spmd_inputs(node, self.rank)
            return self.backend.run_vertex(node, self.rank, ins, VertexEnv(self.context, node.uid, self.rank, 0))
        kind, producer, gb, rest = info
        self._sync()                                           # before the device-side barriers of the aggregate
        uid, (kbits, in_views, key_is_f, _combine) = self._gb_ready
        assert uid == producer.uid
        dec = gb.params["decomp"]
        lanes = dec.gpu_lanes()
        try:
            gk, gl = self._dagg(lanes).run(kbits.contiguous(), [None if c is None else c.contiguous() for c in in_views])
        except OverflowError as ex:
            raise DryadLinqException(ErrorCode.FailureInHashGroupBy, "Arithmetic operation resulted in an overflow.") from ex
        self.backend.fused_aggregates = getattr(self.backend, "fused_aggregates", 0) + 1
        batch = G._group_result_batch(dec, lanes, gk, gl, key_is_f)
        return self._run_rest([batch], node, rest)

    # -- broadcast join ---------------------------------------------------------------------------
    def _run_bjoin(self, node):
        """Join(broadcast=True): this rank's slice of the small side is replicated to every GPU and
        built into each GPU's table by BroadcastJoin (NVLS multicast stores when the box has them,
        else P2P stores); the outer side is probed where it is by the generated kernel."""
        from . import gpu_backend as G
        from ..expr import ir as X
        from ..parallel.djoin import BroadcastJoin
        kind, src, jn, rest = self._fused[node.uid]
        be, ctx = self.backend, self.context
        p = jn.params
        env = VertexEnv(ctx, node.uid, self.rank, 0)
        outer_ports = self.out[node.children[0][0].uid][self.rank]
        outer = be._columnar(outer_ports[0], node.children[0][0].schema)
        inner = be._columnar(self.out[src.uid][self.rank][0], src.schema)
        res = p["result"].ir
        ok = (isinstance(outer, ColumnBatch) and isinstance(inner, ColumnBatch) and not outer.is_rows and not inner.is_rows
              and isinstance(res, (dict, X.Expr)) and p.get("comparer") is None
              and isinstance(p["outer_key"].ir, X.Expr) and isinstance(p["inner_key"].ir, X.Expr)
              and p["outer_key"].ir.kind == p["inner_key"].ir.kind
              and all(v.dim() == 1 for v in inner.cols.values()))
        okb = ikb = None
        if ok:
            okb, ikb = G._key_column(outer, p["outer_key"]), G._key_column(inner, p["inner_key"])
            ok = okb is not None and ikb is not None and not G._has_reserved_key(okb, ikb)
        dev = be.device
        self._sync()
        t = torch.tensor([1 if ok else 0, inner.n if ok else 0], dtype=torch.int64, device=dev)
        sizes = torch.zeros(self.world * 2, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(sizes, t)
        sz = sizes.view(self.world, 2).tolist()
        if not all(f for f, _ in sz):
            return self._bjoin_unfused(node, src, outer_ports)
        total = sum(n for _, n in sz)
        base = sum(n for _, n in sz[:self.rank])
        names = list(inner.cols)
        cache = ctx.__dict__.setdefault("_bjoin_cache", {})
        sig = (len(names),)
        bj = cache.get(sig)
        if bj is None or bj.nrows < total:
            bj = BroadcastJoin(ctx._spmd_arena(), max(total, 1024) if bj is None else 2 * total, len(names))
            cache[sig] = bj
        nrows_saved, bj.nrows = bj.nrows, total               # build over the rows actually present
        batch = None
        try:
            if total == 0:
                raise G.CompileError("empty broadcast side")
            wire = [(v.view(torch.int64) if v.dtype == torch.float64 else v.to(torch.int64)).contiguous() for v in inner.cols.values()]
            bj.build(ikb.contiguous(), wire, base)
            uniq = bool(int((bj.next[:total] >= 0).sum().item()) == 0) if total else True
            if not uniq:
                raise G.CompileError("duplicate keys on the broadcast side")
            proj = res if isinstance(res, dict) else {"value": res}
            hit = X.probe_slot("jt", X.col("__okey"))
            sub, arrays = {}, {}
            for k, v, wcol in zip(names, inner.cols.values(), bj.cols):
                kind_k = "f" if v.dtype == torch.float64 else ("b" if v.dtype == torch.bool else "i")
                g = X.gather("i." + k, hit, "f" if kind_k == "f" else "i")
                sub["i." + k] = X.cast(g, "b") if kind_k == "b" else g
                col = wcol[:max(total, 1)]
                arrays["i." + k] = col.view(torch.float64) if kind_k == "f" else col
            outs = {}
            for name, e in proj.items():
                e = X.wrap(e)
                used = {c for c in e.columns() if c.startswith("i.")}
                outs[name] = e.substitute({c: sub[c] for c in used})
            cols = {("o." + k): v for k, v in outer.cols.items()}
            cols["__okey"] = okb.contiguous()
            need = {a for e in outs.values() for a in arrays if G._mentions_array(e, a)}
            if outer.n:
                out, flags = G.run_on_gpu(outs, cols, outer.n, predicate=hit >= 0, tables={"jt": bj},
                                          arrays={a: bj.slot_payload(arrays[a].contiguous()) for a in need})
                idx, m = G.K.compact_flags(flags)
                batch = ColumnBatch(out, outer.n, tuple_like=G._tuple_cls(res)).take(idx)
            else:
                batch = ColumnBatch({k: torch.empty(0, dtype=self._DT[X.wrap(e).kind], device=dev) for k, e in proj.items()}, 0,
                                    tuple_like=G._tuple_cls(res))
        except (G.CompileError, X.TraceError):
            batch = None
        bj.nrows = nrows_saved
        # every rank must take the same path from here on (the unfused form is a collective)
        done = torch.tensor([0 if batch is None else 1], dtype=torch.int64, device=dev)
        dist.all_reduce(done, op=dist.ReduceOp.MIN)
        if not int(done.item()):
            return self._bjoin_unfused(node, src, outer_ports)
        be.fused_broadcast_joins = getattr(be, "fused_broadcast_joins", 0) + 1
        return self._run_rest([batch], node, rest)

    def _bjoin_unfused(self, node, src, outer_ports):
        """The planned form without pickles: the small side is all-gathered as tensors and the Join
        vertex runs on (outer partition, full small side)."""
        payload = self.out[src.uid][self.rank][0]
        if self._all_columnar([payload], src.schema):
            parts = self._allgather_batches(payload, src.schema, getattr(payload, "tuple_like", False))
            inner_all = [ColumnBatch.concat(parts)]
        else:
            recs = self.backend.to_records(payload)
            everyone: List = [None] * self.world
            dist.all_gather_object(everyone, recs)
            self.object_collectives += 1
            inner_all = [sum(everyone, [])]
        env = VertexEnv(self.context, node.uid, self.rank, 0)
        return self.backend.run_vertex(node, self.rank, [[outer_ports[0]], inner_all], env)

    def _commit_local(self):
        """Output commit, reached only after the graph manager's `commit` command (every stage
        succeeded on every rank).  A table with a store URI is written by all ranks together: each
        rank writes its own partition as an uncommitted version file, the ranks agree that all
        writes succeeded, then every rank renames its part and rank 0 publishes the metadata — a
        failure anywhere removes the version files, so a failed job leaves no part files."""
        import os
        from ..channel import store
        results = {}
        for root in self.plan.roots:
            node = _tail(root)
            mine = self.rank if root.partition_count == self.world else 0
            payload = self.out[root.uid][mine]
            if root.partition_count == self.world:
                # a distributed table: every rank holds ITS partition, the others are None (the same
                # layout from_columns / from_rows build), so a later query sees `world` partitions
                parts = [None] * self.world
                parts[self.rank] = payload[0] if payload is not None else []
            else:
                parts = [payload[0]] if payload is not None else [[]]
            uri = node.params.get("uri")
            if uri and self.world > 1 and not store.supports_part_commit(uri):
                # a provider without per-partition commit: rank 0 collects the partitions and publishes the table
                mine_recs = self.backend.to_records(payload[0]) if payload is not None else []
                gathered: List = [None] * self.world
                dist.all_gather_object(gathered, list(mine_recs))
                self.object_collectives += 1
                if self.rank == 0:
                    n_out = root.partition_count
                    store.write_table(uri, gathered[:n_out] if n_out == self.world else [gathered[0]], root.schema,
                                      node.params.get("compression"))
                dist.barrier()
                table = self.context._commit_table(None, parts, root.schema, root.info, None, node.params.get("temporary", False))
                table.uri = uri
            elif uri and self.world > 1:
                tmp = final = None
                size, err = 0, None
                writes = payload is not None and (root.partition_count == self.world or self.rank == 0)
                try:
                    if writes:
                        tmp, final, size = store.write_part(uri, mine, payload[0], node.params.get("compression"))
                except Exception as ex:  # noqa: BLE001
                    err = ex
                ok = self._cp.agree(err is None) if self._cp is not None else err is None
                if not ok:
                    if tmp and os.path.exists(tmp):
                        os.remove(tmp)
                    raise DryadLinqException(ErrorCode.WriteFileError,
                                             f"output {uri} was not committed: {err or 'another rank could not write its partition'}")
                if writes:
                    store.commit_part(tmp, final)
                nparts = root.partition_count
                sizes: List = [None] * self.world
                dist.all_gather_object(sizes, size)                    # control plane: one integer per rank
                if self.rank == 0:
                    store.finish_table(uri, sizes[:nparts], root.schema, node.params.get("compression"))
                dist.barrier()
                table = self.context._commit_table(None, parts, root.schema, root.info, None, node.params.get("temporary", False))
                table.uri = uri
            else:
                table = self.context._commit_table(uri if self.world == 1 else None, parts, root.schema, root.info,
                                                   node.params.get("compression"), node.params.get("temporary", True))
            results[node.params["qexpr_id"]] = table
        return results
