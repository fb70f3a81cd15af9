"""Distributed k-means iteration (K19 + C3): tcgen05 assign/accumulate kernel per GPU, then an
all-reduce of (sums, counts) and the centroid update — no NCCL.  The all-reduce runs INSIDE the
NVSwitch when the box offers NVLS multicast (multimem.ld_reduce + multimem.st on a symmetric buffer,
parallel/nvls.py); otherwise every rank loads its peers' partials over NVLink in fixed rank order.

Points are sharded by rows (the DP partitioning of the dataset); centroids are replicated.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import _native as N
from ..ops import core as K
from .symm import SymmetricArena

D = 128


class KMeans:
    def __init__(self, arena: SymmetricArena, k: int, n_local_max: int, use_nvls: bool = True, exact: bool = True):
        if k % 128 or k > 2048:
            raise ValueError("k must be a multiple of 128 and at most 2048")
        self.arena, self.k = arena, k
        self.world, self.rank = arena.world, arena.rank
        dev = arena.device
        from .nvls import NvlsBuffer
        self.nvls = NvlsBuffer.create(k * D * 4 + k * 8 + 4096, dev, self.rank, self.world) if use_nvls else None
        if self.nvls is not None:
            self.sums, self.sums_off = self.nvls.alloc((k, D), torch.float32)
            self.counts, self.counts_off = self.nvls.alloc((k,), torch.int64)
            own = lambda t: torch.tensor([t.data_ptr()], dtype=torch.int64).to(dev)   # noqa: E731
            self.peer_sums, self.peer_counts = own(self.sums), own(self.counts)       # totals are local after the all-reduce
        else:
            self.sums, self.sums_off = arena.alloc((k, D), torch.float32)        # partials, peer-readable
            self.counts, self.counts_off = arena.alloc((k,), torch.int64)
            arena.check_symmetric("KMeans partials")
            self.peer_sums = arena.peer_ptr_table(self.sums_off)
            self.peer_counts = arena.peer_ptr_table(self.counts_off)
        self.centroids = torch.empty((k, D), dtype=torch.float32, device=dev)
        self.cnorm = torch.empty(k, dtype=torch.float32, device=dev)
        self.total_counts = torch.zeros(k, dtype=torch.int64, device=dev)
        self.assign = torch.empty(max(n_local_max, 1), dtype=torch.int32, device=dev)
        self.inertia = torch.zeros(1, dtype=torch.float32, device=dev)
        # fp32-exact assignments: tensor-core (TF32) choice certified per point in the assign kernel,
        # uncertified points re-evaluated in fp32 (csrc/kernels/kmeans.cu km_recheck)
        self.exact = exact
        if exact:
            self.sep2 = torch.empty(k, dtype=torch.float32, device=dev)
            self.uncertain = torch.empty(max(n_local_max, 1), dtype=torch.int32, device=dev)
            self.ucount = torch.zeros(2, dtype=torch.int32, device=dev)       # [queued for re-check, changed by it]

    def set_centroids(self, c: torch.Tensor) -> None:
        self.centroids.copy_(c)
        N.check(N.kernels().dryad_kmeans_norms(K.ptr(self.centroids), self.k, K.ptr(self.cnorm), K.stream()), "kmeans_norms")
        K._count()
        self._refresh_sep()

    def _refresh_sep(self) -> None:
        if self.exact:
            N.check(N.kernels().dryad_kmeans_sep(K.ptr(self.centroids), self.k, K.ptr(self.sep2), K.stream()), "kmeans_sep")
            K._count()

    def iterate(self, points: torch.Tensor, max_ctas: int = 0) -> None:
        """One Lloyd iteration on this rank's shard ([n,128] fp32, contiguous).  Leaves the new
        centroids in self.centroids (identical bits on every rank), assignments in self.assign."""
        n = points.shape[0]
        lib = N.kernels()
        if self.world > 1:
            # every rank has finished the previous iteration's update (which reads all peers'
            # partials in the P2P form): only now may the partials be cleared
            self.arena.barrier(slot=2)
        self.sums.zero_()
        self.counts.zero_()
        self.inertia.zero_()
        if self.exact:
            self.ucount.zero_()
            N.check(lib.dryad_kmeans_assign_exact(K.ptr(points), n, K.ptr(self.centroids), K.ptr(self.cnorm), K.ptr(self.sep2),
                                                  self.k, K.ptr(self.assign), K.ptr(self.sums), K.ptr(self.counts),
                                                  K.ptr(self.inertia), K.ptr(self.uncertain), K.ptr(self.ucount),
                                                  self.ucount.data_ptr() + 4, max_ctas, K.stream()), "kmeans_assign_exact")
            K._count(2)
        else:
            N.check(lib.dryad_kmeans_assign(K.ptr(points), n, K.ptr(self.centroids), K.ptr(self.cnorm), self.k,
                                            K.ptr(self.assign), K.ptr(self.sums), K.ptr(self.counts),
                                            K.ptr(self.inertia), max_ctas, K.stream()), "kmeans_assign")
            K._count()
        if self.world > 1:
            self.arena.barrier(slot=3)         # every rank's partials are complete and visible
        if self.nvls is not None:
            # in-switch all-reduce, two-shot: every copy ends up with identical totals
            self.nvls.allreduce_f32(self.sums_off, self.k * D)
            self.nvls.allreduce_u64(self.counts_off, self.k)
            self.arena.barrier(slot=2)         # all slices have been multicast back
        N.check(lib.dryad_kmeans_update(K.ptr(self.peer_sums), K.ptr(self.peer_counts), 1 if self.nvls is not None else self.world, self.k,
                                        K.ptr(self.centroids), K.ptr(self.cnorm), K.ptr(self.total_counts),
                                        K.stream()), "kmeans_update")
        K._count()
        self._refresh_sep()

    def iterate_graph(self, points: torch.Tensor) -> None:
        """Single-GPU iterations as a CUDA graph (sub-graph vertex, runtime/graph_vertex.py): the five
        launches of an iteration (two clears, assign, update, inertia clear) replay as one.  The
        capture is keyed by the point buffer; multi-GPU iterations carry epoch-numbered barriers in
        their kernel parameters and run eagerly."""
        if self.world > 1:
            return self.iterate(points)
        key = (points.data_ptr(), points.shape[0])
        if getattr(self, "_graph_key", None) != key:
            from ..runtime.graph_vertex import GraphVertex
            self._graph = GraphVertex(lambda: self.iterate(points), device=points.device).capture()
            self._graph_key = key
        self._graph.run()
        K._count(2)
