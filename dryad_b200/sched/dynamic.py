"""Run-time graph rewriting: the dynamic managers of the graph-manager process.

Reference: GraphManager/stagemanager/DrDynamicAggregateManager.h:24-179 (aggregation tree: as the
producers of an N:1 edge finish, outputs that sit on the same machine — then pod, then anywhere — are
folded by internal vertices before the consumer sees them), DrDynamicBroadcast.cpp:125-130 (a producer
with many consumers gets floor(sqrt(n)) copy vertices between itself and them), and
DrDynamicDistributor.h:25-124 / DrDynamicRangeDistributor.h:22-78 (the number of consumers of a
repartitioning stage follows the data volume).  The planner only TAGS stages
(LinqToDryad/DynamicManager.cs:35-202 -> `DLinqQueryNode.dynamic_manager`); everything here happens
while the job runs, driven by vertex completions, through `GraphManager.splice` /
`GraphManager.resize_stage` of the native graph.

The managers are owned by `sched.jm.JobManager` and are called from its pump handler only, so they
see and change the graph single-threaded.
"""
from __future__ import annotations

import json
from typing import Dict, List, Optional, Tuple

from ..plan.qnodes import DynamicManagerType, NodeType

from ..runtime.vertex_host import DYN_UID_BASE  # noqa: E402


class DynNode:
    """Stands for the plan node of an internal / copy vertex (they have no node in the client's plan)."""

    def __init__(self, uid: int, name: str, kind: str, consumer=None, port: int = 0):
        self.uid, self.name, self.kind, self.consumer, self.port = uid, name, kind, consumer, port
        self.partition_count = 0
        self.children: list = []

    class _T:
        value = "Dynamic"
    type = _T()


class DynamicManagers:
    def __init__(self, jm):
        self.jm = jm
        ctx = jm.ctx
        self.fan_in = int(getattr(ctx, "aggregation_tree_fan_in", 0) or 0)          # 0 = aggregation tree off
        if self.fan_in and jm.app.max_aggregate_inputs != type(jm.app)().max_aggregate_inputs:
            self.fan_in = jm.app.max_aggregate_inputs                               # -mai on the manager's command line wins
        self.broadcast_min = int(getattr(ctx, "broadcast_tree_min_consumers", 0) or 0)   # 0 = copy vertices off
        self.data_per_vertex = int(getattr(ctx, "dynamic_data_per_vertex", 0) or 0)      # 0 = distributor off
        self.agg: Dict[int, dict] = {}              # consumer gm vertex -> state of its tree
        self.stage_of: Dict[Tuple[str, int], int] = {}      # (kind, consumer/producer stage uid) -> dynamic gm stage
        self.subst: Dict[int, Dict[str, Optional[str]]] = {}   # consumer gm vertex -> {"U/s": "X/p" | None}
        self.spliced = {"aggregate": 0, "copy": 0, "resized": 0}
        self._bcast_done: set = set()
        self.dist: Dict[int, dict] = {}             # distributor stage uid -> state
        self.widths: Dict[int, int] = {}            # stage uid -> partition count decided at run time
        self.nports: Dict[int, int] = {}            # distributor stage uid -> output ports decided at run time
        self.pcount: Dict[int, int] = {}            # range-sampler stage uid -> number of ranges decided at run time
        self._hold_by_uid: Dict[int, int] = {}      # stage whose vertices are held -> distributor stage uid
        self._find_candidates()
        self._find_distributors()

    # ---- which vertices are eligible --------------------------------------------------------------
    def _find_candidates(self) -> None:
        from ..runtime.vertex_ops import combiner_of
        jm = self.jm
        if not self.fan_in:
            return
        from ..plan.qnodes import ConnOp
        agg_tags = (DynamicManagerType.FullAggregator, DynamicManagerType.PartialAggregator)
        for node in jm.plan.nodes:
            if len(node.children) != 1:
                continue
            child, op = node.children[0]
            # a stage's tag is its tail's; the partial aggregate sits before the HashPartition in its super node
            tagged = child.dynamic_manager in agg_tags or any(m.dynamic_manager in agg_tags for m in (child.members or []))
            if child.partition_count <= self.fan_in or not (tagged or getattr(jm.ctx, "aggregation_tree_all", False)):
                continue
            cross = op == ConnOp.CrossProduct
            if not cross and node.partition_count != 1:
                continue                                  # N:1 gather, or N:M where consumer j reads port j of every producer
            found = combiner_of(node)
            if found is None:
                continue
            for j in range(node.partition_count):
                v = jm.vid[(node.uid, j)]
                self.agg[v] = {"node": node, "part": j, "port": j if cross else 0, "commutative": found[1], "done": False,
                               "level": 0}

    def _dyn_stage(self, kind: str, key_uid: int, name: str, consumer=None, port: int = 0) -> int:
        jm = self.jm
        k = (kind, key_uid, port)
        st = self.stage_of.get(k)
        if st is None:
            st = jm.gm.add_stage(name, 0)
            self.stage_of[k] = st
            jm.nodes[st] = DynNode(DYN_UID_BASE + st, name, kind, consumer, port)
        return st

    # ---- aggregation tree ---------------------------------------------------------------------------
    def _splice_group(self, cv: int, state: dict, members: List[int]) -> None:
        """One internal vertex over `members` (gm vertex ids, inputs of consumer `cv`)."""
        jm = self.jm
        node = state["node"]
        st = self._dyn_stage("combine", node.uid, f"Agg__{node.name}", node, state["port"])
        dyn = jm.nodes[st]
        try:
            x = jm.gm.splice(st, members, [cv])
        except ValueError:
            state["done"] = True                     # the consumer belongs to a start clique or has started: leave it alone
            return
        xp = jm.gm.vertex(x)["part"]
        dyn.partition_count = xp + 1
        jm.vid[(dyn.uid, xp)] = x
        jm.vkey[x] = (dyn.uid, xp)
        ins = jm.ex._gm_inputs.setdefault(cv, [])
        sub = self.subst.setdefault(cv, {})
        keys = [jm.vkey[m] for m in members]
        first = min(range(len(keys)), key=lambda i: ins.index(keys[i]))
        for i, key in enumerate(keys):
            ins.remove(key)
            # in the consumer's port order the group's result stands where its first member stood
            sub[f"{key[0]}/{key[1]}"] = f"{dyn.uid}/{xp}" if i == first else None
        ins.append((dyn.uid, xp))
        jm.ex._gm_inputs[x] = list(keys)
        state["grouped"] = state.get("grouped", set()) | set(members)
        self.spliced["aggregate"] += 1
        jm.reporter._emit("dynamic", manager="aggregate", consumer=f"{node.uid}.{state['part']}", vertex=f"{dyn.uid}.{xp}",
                          inputs=" ".join(f"{a}.{b}" for a, b in keys), level=state["level"])

    def on_completed(self, v: int, host: int) -> None:
        """Called after vertex v's completion was applied to the graph and before anything is dispatched."""
        self._aggregate(v, host)
        self._broadcast(v)
        self._distribute()

    # ---- dynamic distributor ------------------------------------------------------------------------
    def _find_distributors(self) -> None:
        from ..plan.qnodes import ConnOp
        from ..runtime.executor import JobExecutor
        jm = self.jm
        if not self.data_per_vertex:
            return
        xp = JobExecutor.edge_is_cross_product
        for node in jm.plan.nodes:
            tail = node.members[-1] if node.type == NodeType.Super else node
            is_hash = tail.type == NodeType.HashPartition and node.dynamic_manager == DynamicManagerType.HashDistributor
            is_range = (tail.type == NodeType.RangePartition and node.dynamic_manager == DynamicManagerType.RangeDistributor
                        and tail.params.get("range_keys") is None and node.type != NodeType.Super and len(node.children) == 2)
            if not (is_hash or is_range):
                continue
            if not node.children or len(node.parents) != 1:
                continue
            sampler = None
            if is_range:
                # the separators come from a one-vertex stage that must produce (new width - 1) of them: IT is held
                sampler = node.children[1][0]
                stail = sampler.members[-1] if sampler.type == NodeType.Super else sampler
                if stail.type != NodeType.RangeSampler or sampler.partition_count != 1:
                    continue
            cons = node.parents[0]
            op = next(o for c, o in cons.children if c is node)
            if op != ConnOp.CrossProduct or cons.partition_count != tail.params.get("nports") or len(cons.children) != 1:
                continue
            # the stages that change width with the consumer (pointwise, same width), as the native resize follows them
            old = cons.partition_count
            follow = [cons]
            ok = True
            special = (NodeType.Concat, NodeType.Zip, NodeType.Reverse, NodeType.SlidingWindow)
            for n in follow:
                for par in n.parents:
                    pop = next(o for c, o in par.children if c is n)
                    phead = par.members[0].type if par.type == NodeType.Super else par.type
                    if phead in special:
                        ok = False                                    # these read neighbouring partitions by planned position
                        continue
                    if xp(par, n, pop):
                        continue                                      # cross product: any widths connect
                    if par.partition_count == old:
                        if par not in follow:
                            follow.append(par)
                    elif par.partition_count != 1:
                        ok = False                                    # grouped pointwise edge: widths must divide
            for n in follow[1:]:
                for c, o in n.children:
                    if c not in follow and not xp(n, c, o) and c.partition_count != 1:
                        ok = False                                    # a second same-width input (co-partitioned join)
            if ok:
                data = [c for c, _ in node.children if c is not sampler]
                self.dist[node.uid] = {"node": node, "tail": tail, "consumer": cons, "follow": follow, "decided": False,
                                       "held": [], "hold_uid": sampler.uid if sampler is not None else node.uid,
                                       "data": data, "sampler": sampler}
                self._hold_by_uid[self.dist[node.uid]["hold_uid"]] = node.uid

    def hold(self, v: int, version: int) -> bool:
        """A distributor vertex is not dispatched before its stage's output width has been decided."""
        key = self.jm.vkey.get(v)
        st = self.dist.get(self._hold_by_uid.get(key[0], -1)) if key else None
        if st is None or st["decided"]:
            return False
        st["held"].append((v, version))
        return True

    def _distribute(self) -> None:
        from .gm import consumer_count
        jm = self.jm
        for uid, st in self.dist.items():
            if st["decided"]:
                continue
            node = st["node"]
            ups = [(c.uid, i) for c in st["data"] for i in range(c.partition_count)]
            if any(k not in jm.versions for k in ups):
                continue
            st["decided"] = True
            total = sum(jm.out_bytes.get(k, 0) for k in ups)
            old = st["consumer"].partition_count
            new = consumer_count(total, self.data_per_vertex, 1, 20000)
            if new != old:
                stage = jm.stage_of_uid[st["consumer"].uid]
                n = jm.gm.resize_stage(stage, new, follow=True)
                for f in st["follow"]:
                    f.partition_count = new
                    self.widths[f.uid] = new
                st["tail"].params["nports"] = new
                node.n_out_ports = new
                self.nports[uid] = new
                if st["sampler"] is not None:
                    smp = st["sampler"]
                    (smp.members[-1] if smp.type == NodeType.Super else smp).params["pcount"] = new
                    self.pcount[smp.uid] = new
                jm.rebuild_maps()
                self.spliced["resized"] += n
                jm.reporter._emit("dynamic", manager="distributor", stage=uid, bytes=total, consumers_before=old,
                                  consumers=new, stages_resized=n)
            held, st["held"] = st["held"], []
            for v, version in held:
                jm.enqueue([(v, version)])

    def _aggregate(self, v: int, host: int) -> None:
        jm = self.jm
        key = jm.vkey.get(v)
        for cv, state in self.agg.items():
            if state["done"] or key not in jm.ex._gm_inputs.get(cv, []):
                continue
            info = jm.gm.vertex(cv)
            if info["state"] not in ("Waiting", "Ready"):
                state["done"] = True
                continue
            ins = [jm.vid[k] for k in jm.ex._gm_inputs[cv]]
            done = [i for i in ins if jm.vkey[i] in jm.versions]
            grouped = state.get("grouped", set())
            if state["commutative"] and len(done) < len(ins):
                # incremental, by machine: a full group on one host is folded while the others still run
                here = [i for i in done if i not in grouped and jm.where.get(jm.vkey[i]) == host]
                if len(here) >= self.fan_in:
                    self._splice_group(cv, state, here[:self.fan_in])
                continue
            if len(done) < len(ins):
                continue
            # every current input is there: group what is left (machine, then anywhere), or finish
            if len(ins) <= self.fan_in:
                state["done"] = True
                continue
            state["level"] += 1
            if state["commutative"]:
                sizes = [max(1, jm.out_bytes.get(jm.vkey[i], 0)) for i in ins]
                chunks = []
                for by_machine in (True, False):     # same machine first, then anywhere (one pod on this box)
                    machine = [jm.where.get(jm.vkey[i], -1) if by_machine else 0 for i in ins]
                    jm.app.max_aggregate_inputs = self.fan_in
                    gids = jm.app.aggregate_groups(sizes, machine, [0] * len(ins))     # -at / -nca apply here
                    groups: Dict[int, List[int]] = {}
                    for i, g in zip(ins, gids):
                        groups.setdefault(g, []).append(i)
                    chunks = [m for m in groups.values() if len(m) >= 2]
                    if chunks:
                        break
            else:
                order = list(ins)                    # contiguous runs keep the port order
                chunks = [order[a:a + self.fan_in] for a in range(0, len(order), self.fan_in)]
                chunks = [c for c in chunks if len(c) >= 2]
            if not chunks:
                state["done"] = True
                continue
            for members in chunks:
                self._splice_group(cv, state, members)

    # ---- broadcast tree -----------------------------------------------------------------------------
    def _broadcast(self, v: int) -> None:
        from .gm import broadcast_tree
        jm = self.jm
        if not self.broadcast_min or v in self._bcast_done:
            return
        key = jm.vkey.get(v)
        if key is None or key[0] >= DYN_UID_BASE:
            return
        node = jm.node_by_uid[key[0]]
        if node.partition_count != 1:
            return
        self._bcast_done.add(v)
        from ..plan.qnodes import ConnOp
        for parent in node.parents:
            if parent.partition_count < max(5, self.broadcast_min):
                continue
            head = parent.members[0].type if parent.type == NodeType.Super else parent.type
            edge = next(((c, op) for c, op in parent.children if c is node), None)
            if edge is None:
                continue
            # a tagged Tee replicates: its ports are identical, so a copy of port 0 serves every consumer;
            # untagged producers qualify only over plain 1:N edges (every consumer reads port 0)
            tagged = node.dynamic_manager == DynamicManagerType.Broadcast
            plain = edge[1] != ConnOp.CrossProduct and head not in (NodeType.ForkBranch, NodeType.Concat, NodeType.Zip,
                                                                    NodeType.SlidingWindow, NodeType.Reverse)
            if not (tagged or (plain and getattr(jm.ctx, "broadcast_tree_all", False))):
                continue
            cons = [jm.vid[(parent.uid, i)] for i in range(parent.partition_count)]
            cons = [c for c in cons if jm.gm.vertex(c)["state"] in ("Waiting", "Ready") and key in jm.ex._gm_inputs.get(c, [])]
            copies, copy_of = broadcast_tree(len(cons))
            if copies <= 0:
                continue
            st = self._dyn_stage("copy", node.uid, f"CP__{node.name}", None, 0)
            dyn = jm.nodes[st]
            for cpy in range(copies):
                mine = [c for c, which in zip(cons, copy_of) if which == cpy]
                if not mine:
                    continue
                try:
                    x = jm.gm.splice(st, [v], mine)
                except ValueError:
                    continue                         # a consumer in a start clique: it keeps reading the producer directly
                xp = jm.gm.vertex(x)["part"]
                dyn.partition_count = xp + 1
                jm.vid[(dyn.uid, xp)] = x
                jm.vkey[x] = (dyn.uid, xp)
                jm.ex._gm_inputs[x] = [key]
                for c in mine:
                    ins = jm.ex._gm_inputs[c]
                    ins.remove(key)
                    ins.append((dyn.uid, xp))
                    self.subst.setdefault(c, {})[f"{key[0]}/{key[1]}"] = f"{dyn.uid}/{xp}"
                self.spliced["copy"] += 1
                jm.reporter._emit("dynamic", manager="broadcast", producer=f"{key[0]}.{key[1]}", vertex=f"{dyn.uid}.{xp}",
                                  consumers=len(mine))

    # ---- what the vertex host needs to know ---------------------------------------------------------
    def command_args(self, v: int) -> List[str]:
        jm = self.jm
        node = jm.nodes[jm.gm.vertex(v)["stage"]]
        out: List[str] = []
        if isinstance(node, DynNode):
            out += ["internal", node.kind, str(node.consumer.uid if node.consumer is not None else -1), str(node.port)]
        sub = self.subst.get(v)
        if sub:
            out.append("subst=" + json.dumps(sub))
        if self.widths or self.nports:
            out.append("widths=" + json.dumps({"widths": self.widths, "nports": self.nports, "pcount": self.pcount}))
        return out
