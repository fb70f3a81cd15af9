#!/usr/bin/env python
"""Headline benchmark (BASELINE.json): TeraSort-style RangePartition + OrderBy over 100-byte
synthetic records on N B200s of one node; GroupBy-aggregate rows/s is reported alongside.

  python bench.py --gpus N --steps K --warmup W          (N>1: launched under torchrun)
  python bench.py --impl reference ...                   (reference arm: unavailable)

`value` = whole-job sorted GB/s (N * rows_per_gpu * 100 bytes / step time), device-timed with CUDA
events, max over ranks, weak scaling (rows per GPU fixed).  The timed step goes from unsorted rows
resident in HBM to globally sorted rows resident in HBM: sampler, separators, fused range shuffle
over NVLink, radix sort, tie repair, payload gather.  `e2e` runs the same sort through the public
query API with the host->device copy of the rows (pinned memory) and a device->host read of the
validation summary inside every timed step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "terasort_sorted_GB_per_s"


def _clock_sampler(gpu_index: int, stop: threading.Event, out: list):
    q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")
    try:
        p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100",
                              "-i", str(gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    except Exception:  # noqa: BLE001
        return
    try:
        while not stop.is_set():
            line = p.stdout.readline()
            if not line:
                break
            out.append(line.strip())
    finally:
        p.kill()


def _clock_summary(lines):
    sm, mx, reasons = [], 0.0, set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for l in lines:
        f = [x.strip() for x in l.split(",")]
        if len(f) < 6:
            continue
        try:
            sm.append(float(f[0]))
            mx = max(mx, float(f[1]))
        except ValueError:
            continue
        for nm, v in zip(names, f[2:6]):
            if v.lower().startswith("active"):
                reasons.add(nm)
    if not sm:
        return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
    busy = [x for x in sm if x > 0.6 * max(sm)] or sm
    return {"sm_mhz": statistics.median(busy), "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--rows-per-gpu", type=int, default=40_000_000)
    ap.add_argument("--groupby-rows-per-gpu", type=int, default=125_000_000)
    ap.add_argument("--groupby-keys", type=int, default=1 << 20)
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--skip-groupby", action="store_true")
    ap.add_argument("--kmeans-points", type=int, default=125_000_000, help="points per GPU of the named-size k-means run (0 = skip)")
    ap.add_argument("--skip-baseline", action="store_true", help="do not time the PyTorch+NCCL comparison arm")
    a = ap.parse_args()

    if a.impl == "reference":
        print(json.dumps({"impl": "reference", "unavailable":
                          "Liuqin/Dryad is a C#/.NET 4.5 + C++/CLI Visual Studio 2013 solution for Windows/YARN; "
                          "pip cannot install it (no setup.py/pyproject) and no .NET toolchain exists in this image"}))
        return 0

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != a.gpus:
        if world == 1 and a.gpus > 1:
            print(f"bench.py: --gpus {a.gpus} needs torchrun (WORLD_SIZE={world})", file=sys.stderr)
            return 2
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    from dryad_b200.utils.numa import bind_to_gpu
    bound = bind_to_gpu(local_rank)            # pinned staging memory is first-touched on the GPU's socket
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    from dryad_b200 import DryadLinqContext, RowKey
    from dryad_b200.models import terasort
    from dryad_b200.ops import core as K
    from dryad_b200.parallel.dsort import RowSorter
    from dryad_b200.parallel.symm import SymmetricArena

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n = a.rows_per_gpu
    W = terasort.RECORD_BYTES
    rows = terasort.generate_rows(n, seed=7, device=dev, rank=rank)

    # ---- kernel-path step: HBM-resident rows -> globally sorted HBM-resident rows ----------------
    arena_bytes = (int(n * 1.12) + 8192) * (W + 8) + (64 << 20) if world > 1 else (1 << 20)
    arena = SymmetricArena(arena_bytes, rank, world, None, dev)
    sorter = RowSorter(arena, n, W, key_off=0, key_len=terasort.KEY_BYTES)

    def step():
        return sorter.sort(rows)

    for _ in range(a.warmup):
        step()
    barrier()
    samples, stop = [], threading.Event()
    th = threading.Thread(target=_clock_sampler, args=(local_rank, stop, samples), daemon=True)
    th.start()
    l0 = K.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(a.steps):
        out, m = step()
    e1.record()
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    launches = K.launch_count() - l0
    stop.set()
    th.join(timeout=2)
    clocks = _clock_summary(samples)
    ms_step = ms_total / a.steps
    gbs = world * n * W / (ms_step * 1e-3) / 1e9

    # correctness of the last step: every rank's slice is sorted, the multiset of keys is preserved
    chk = K.check_sorted_rows(out, 0, terasort.KEY_BYTES).cpu()
    viol = int(chk[0])
    cs_out = int(chk[1]) & ((1 << 64) - 1)
    cs_in = int(K.check_sorted_rows(rows, 0, terasort.KEY_BYTES).cpu()[1]) & ((1 << 64) - 1)
    tot = torch.tensor([viol, m, cs_out - (1 << 63) if cs_out >= (1 << 63) else cs_out,
                        cs_in - (1 << 63) if cs_in >= (1 << 63) else cs_in], dtype=torch.int64, device=dev)
    if world > 1:
        gathered = [torch.zeros_like(tot) for _ in range(world)]
        dist.all_gather(gathered, tot)
    else:
        gathered = [tot]
    g = torch.stack(gathered).cpu()
    ok = int(g[:, 0].sum()) == 0 and int(g[:, 1].sum()) == world * n and \
        (int(g[:, 2].sum()) - int(g[:, 3].sum())) % (1 << 64) == 0

    def timed(fn, steps, warmup):
        """ms per call: device-timed with CUDA events, barrier + synchronize on both sides, max over ranks."""
        for _ in range(warmup):
            fn()
        barrier()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(steps):
            fn()
        t1.record()
        barrier()
        return max_over_ranks(t0.elapsed_time(t1)) / steps

    # ---- the comparison arm the bench line carries itself: PyTorch + NCCL (+cuBLAS) on the same box ---
    # (dryad_b200/baseline/torch_nccl.py: all_to_all_single / torch.sort / unique / index_add_ /
    # all_gather / all_reduce / matmul; it calls none of this repo's kernels)
    from dryad_b200.baseline import torch_nccl as B
    base = {"impl": "pytorch+nccl(+cublas), same process, same box, same data", "steps": None}
    bsteps, bwarm = max(2, min(a.steps, 3)), 1
    base["steps"], base["warmup"] = bsteps, bwarm
    if not a.skip_baseline:
        try:
            del out
            b_ms = timed(lambda: B.terasort_step(rows, world, rank), bsteps, bwarm)
            base["terasort"] = {"ms_per_step": b_ms, "GB_per_s": world * n * W / (b_ms * 1e-3) / 1e9}
        except Exception as ex:  # noqa: BLE001
            base["terasort"] = {"error": f"{type(ex).__name__}: {ex}"}
        torch.cuda.empty_cache()

    # ---- end to end through the public API -------------------------------------------------------
    e2e = None
    if not a.skip_e2e:
        try:
            host_rows = torch.empty((n, W), dtype=torch.uint8).pin_memory()
            host_rows.copy_(rows)
            del sorter, arena
            torch.cuda.empty_cache()
            ctx = DryadLinqContext(world, platform="B200" if world > 1 else "LOCAL", device=f"cuda:{local_rank}")

            def e2e_step():
                q = ctx.from_rows(host_rows, key_off=0, key_len=terasort.KEY_BYTES).order_by(RowKey(0, terasort.KEY_BYTES))
                res = q.submit_and_wait()
                return terasort.validate(res)                 # D2H: [violations, checksum, count]

            # context for the reader: this box's raw pinned-host -> device rate (not part of the timed region)
            probe = torch.empty_like(host_rows, device=dev)
            p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            probe.copy_(host_rows, non_blocking=True)
            p0.record()
            probe.copy_(host_rows, non_blocking=True)
            p1.record()
            torch.cuda.synchronize()
            h2d_rate = n * W / (p0.elapsed_time(p1) * 1e-3) / 1e9
            del probe
            for _ in range(max(1, min(a.warmup, 3))):
                summ = e2e_step()
            barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            for _ in range(a.steps):
                summ = e2e_step()
            s1.record()
            barrier()
            e_ms = max_over_ranks(s0.elapsed_time(s1)) / a.steps
            e2e = {"value": world * n * W / (e_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": e_ms,
                   "what": "load + sort: rows H2D from pinned memory, sorted rows stay in HBM, a 24-byte validation summary comes back",
                   "h2d_bytes_per_step": n * W, "d2h_bytes_per_step": 24, "cpus_bound": len(bound) if bound else 0, "box_h2d_GBps_per_gpu": round(h2d_rate, 1),
                   "violations": int(summ[0]), "rows_out_local": int(summ[2])}
            # variant that also stores: the sorted rows are copied back to pinned host memory every step
            host_out = torch.empty((int(n * 1.13) + 8192, W), dtype=torch.uint8).pin_memory()

            def e2e_rows_back():
                q = ctx.from_rows(host_rows, key_off=0, key_len=terasort.KEY_BYTES).order_by(RowKey(0, terasort.KEY_BYTES))
                res = q.submit_and_wait()
                r_ = terasort.result_rows(res)
                host_out[:r_.shape[0]].copy_(r_, non_blocking=True)
                torch.cuda.current_stream().synchronize()
                return r_.shape[0]

            rb_ms = timed(e2e_rows_back, max(2, a.steps // 2), 1)
            e2e["rows_back"] = {"value": world * n * W / (rb_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_step": rb_ms,
                                "what": "load + sort + store: the sorted rows are also copied to pinned host memory",
                                "h2d_bytes_per_step": n * W, "d2h_bytes_per_step": n * W}
            if not a.skip_baseline:
                def base_e2e():
                    r_ = host_rows.to(dev, non_blocking=True)
                    o_ = B.terasort_step(r_, world, rank)
                    return o_[0, :8].cpu()                      # D2H: 8 bytes
                be_ms = timed(base_e2e, bsteps, bwarm)
                base["terasort_e2e"] = {"ms_per_step": be_ms, "GB_per_s": world * n * W / (be_ms * 1e-3) / 1e9,
                                        "h2d_bytes_per_step": n * W, "d2h_bytes_per_step": 8}
            del host_out
            ctx.dispose()                        # frees this context's symmetric arena (a third of HBM) for what follows
        except Exception as ex:  # noqa: BLE001
            e2e = {"error": f"{type(ex).__name__}: {ex}"}

    # ---- second headline metric: GroupBy-aggregate rows/s -----------------------------------------
    groupby = None
    if not a.skip_groupby:
        try:
            from dryad_b200.models import groupby as gbm
            # headline: general (sparse 64-bit) keys on the hashed table; the dense-id case, where the
            # table is direct-addressed, is the second field
            groupby = gbm.bench(rank, world, dev, a.groupby_rows_per_gpu, a.groupby_keys, steps=a.steps,
                                warmup=a.warmup, max_over_ranks=max_over_ranks, barrier=barrier, sparse=True)
            torch.cuda.empty_cache()
            groupby["dense_ids"] = gbm.bench(rank, world, dev, a.groupby_rows_per_gpu, a.groupby_keys, steps=a.steps,
                                             warmup=a.warmup, max_over_ranks=max_over_ranks, barrier=barrier, sparse=False)
            torch.cuda.empty_cache()
            if not a.skip_baseline:
                g_n = a.groupby_rows_per_gpu
                gk_, gv_ = gbm.generate(g_n, a.groupby_keys, 11, dev, rank, sparse=True)
                ms_ = timed(lambda: B.groupby_hashed_step(gk_, gv_, world), bsteps, bwarm)
                base["groupby_sparse_keys"] = {"ms_per_step": ms_, "rows_per_s": world * g_n / (ms_ * 1e-3),
                                               "how": "unique + index_add_ partials, all_to_all_single by key hash, unique + index_add_"}
                del gk_, gv_
                gk_, gv_ = gbm.generate(g_n, a.groupby_keys, 11, dev, rank, sparse=False)
                ms_ = timed(lambda: B.groupby_step(gk_, gv_, a.groupby_keys, world), bsteps, bwarm)
                base["groupby_dense_ids"] = {"ms_per_step": ms_, "rows_per_s": world * g_n / (ms_ * 1e-3),
                                             "how": "index_add_ / bincount on the dense id + all_reduce"}
                del gk_, gv_
                torch.cuda.empty_cache()
        except Exception as ex:  # noqa: BLE001
            groupby = {"error": f"{type(ex).__name__}: {ex}"} if groupby is None else dict(groupby, baseline_error=f"{type(ex).__name__}: {ex}")
        if not a.skip_e2e and isinstance(groupby, dict) and "error" not in groupby:
            try:
                torch.cuda.empty_cache()
                groupby["e2e"] = gbm.bench_api(rank, world, dev, a.groupby_rows_per_gpu, a.groupby_keys, steps=max(2, a.steps // 2),
                                               warmup=2, max_over_ranks=max_over_ranks, barrier=barrier)
            except Exception as ex:  # noqa: BLE001
                groupby["e2e"] = {"error": f"{type(ex).__name__}: {ex}"}

    # ---- further BASELINE.json configs, short runs: broadcast join (4) and k-means (5) -------------
    extra = {}
    if not a.skip_groupby:
        try:
            from dryad_b200.models import joins
            extra["broadcast_join"] = joins.bench(rank, world, dev, 50_000_000, 1 << 20, steps=a.steps, warmup=a.warmup,
                                                  max_over_ranks=max_over_ranks, barrier=barrier)
            if not a.skip_baseline:
                torch.cuda.empty_cache()
                fk_, fv_, dk_, da_, _ = joins.generate(50_000_000, 1 << 20, 13, dev, rank, world)
                ms_ = timed(lambda: B.join_step(fk_, fv_, dk_, da_, 1 << 20, world), bsteps, bwarm)
                base["broadcast_join"] = {"ms_per_step": ms_, "rows_per_s": world * 50_000_000 / (ms_ * 1e-3),
                                          "how": "all_gather of the dimension table, dense table[key] gather, masked multiply, sum"}
                del fk_, fv_, dk_, da_
            extra["shuffle_hash_join"] = joins.bench_shuffle_join(rank, world, dev, 20_000_000, 4_000_000, steps=max(2, a.steps // 2),
                                                                  warmup=1, max_over_ranks=max_over_ranks, barrier=barrier)
        except Exception as ex:  # noqa: BLE001
            extra.setdefault("broadcast_join", {"error": f"{type(ex).__name__}: {ex}"})
            extra.setdefault("shuffle_hash_join", {"error": f"{type(ex).__name__}: {ex}"})
        try:
            from dryad_b200.models import kmeans as KM
            from dryad_b200.parallel.kmeans import KMeans
            torch.cuda.empty_cache()
            npts, kk = 8_000_000, 1024
            pts, centres = KM.generate(npts, kk, 5, dev, rank)
            karena = SymmetricArena(16 << 20, rank, world, None, dev)
            km = KMeans(karena, kk, npts)
            km.set_centroids(centres + 0.05)
            for _ in range(2):
                km.iterate(pts)
            barrier()
            k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            k0.record()
            for _ in range(a.steps):
                km.iterate(pts)
            k1.record()
            barrier()
            kms = max_over_ranks(k0.elapsed_time(k1)) / a.steps
            queued, changed = km.ucount.tolist()
            extra["kmeans"] = {"metric": "kmeans_seconds_per_iteration", "value": kms * 1e-3, "unit": "s",
                               "points_per_gpu": npts, "dims": 128, "k": kk, "points_per_s": world * npts / (kms * 1e-3),
                               "tf32_tflops_per_gpu": 2.0 * npts * kk * 128 / (kms * 1e-3) / 1e12,
                               "assignments": "fp32-exact: TF32 tensor-core choice certified per point in the assign kernel, "
                                              "uncertified points re-evaluated in fp32",
                               "uncertified_points_last_iteration": int(queued), "moved_by_recheck": int(changed),
                               "total_count_ok": int(km.total_counts.sum().item()) == world * npts}
            # the same iteration without the certificate (TF32 choice taken as is), for the cost of exactness
            km_t = KMeans(karena, kk, npts, exact=False)
            km_t.set_centroids(centres + 0.05)
            extra["kmeans"]["tf32_only_ms_per_iteration"] = timed(lambda: km_t.iterate(pts), a.steps, 2)
            del km_t
            # ... and as a query: do_while over apply_per_partition(KMeansStep) through the public API
            try:
                kctx = DryadLinqContext(world, platform="B200" if world > 1 else "LOCAL", device=f"cuda:{local_rank}")
                kstep = KM.KMeansStep(kctx)
                pq, cq = kctx.from_columns({"x": pts}), kctx.from_columns({"c": centres + 0.05})

                class _Budget:
                    device = True

                    def __init__(self):
                        self.left = 0

                    def __call__(self, new_c, old_c):
                        self.left -= 1
                        return {"value": torch.tensor([self.left > 0], device=dev)}
                bud = _Budget()

                def api_loop(iters):
                    bud.left = iters
                    return cq.do_while(lambda cur: pq.apply_per_partition(kstep, cur),
                                       lambda cur, nxt: nxt.apply_per_partition(bud, cur))
                api_loop(2)
                barrier()
                t_api = time.perf_counter()
                api_loop(5)
                barrier()
                extra["kmeans"]["api_do_while_s_per_iteration"] = max_over_ranks((time.perf_counter() - t_api) / 5)
                kctx.dispose()
            except Exception as ex:  # noqa: BLE001
                extra["kmeans"]["api_do_while_error"] = f"{type(ex).__name__}: {ex}"
            if not a.skip_baseline:
                c0 = centres + 0.05
                ms_ = timed(lambda: B.kmeans_step(pts, c0, world), 2, 1)
                base["kmeans"] = {"ms_per_iteration": ms_, "how": "cuBLAS TF32 matmul + argmin + index_add_ + all_reduce"}
                ms_ = timed(lambda: B.gemm_tf32_only(pts, c0), 3, 1)
                base["kmeans_gemm_only_cublas_tf32"] = {"ms": ms_, "tflops": 2.0 * npts * kk * 128 / (ms_ * 1e-3) / 1e12,
                                                        "what": "the distance GEMM alone (8 M x 128 x 1024), no argmin, no sums, no all-reduce"}
            # BASELINE.json config 5 at its named size: 1 B points x 128 over 8 GPUs = 125 M points per GPU
            # (64 GB of fp32 points per GPU); same per-GPU size at every N (weak scaling)
            if a.kmeans_points > npts:
                try:
                    del pts, km
                    torch.cuda.empty_cache()
                    bigp, centres = KM.generate(a.kmeans_points, kk, 5, dev, rank)
                    kmb = KMeans(karena, kk, a.kmeans_points)
                    kmb.set_centroids(centres + 0.05)
                    bms = timed(lambda: kmb.iterate(bigp), 3, 1)
                    extra["kmeans_named_size"] = {"metric": "kmeans_seconds_per_iteration", "value": bms * 1e-3, "unit": "s",
                                                  "points_per_gpu": a.kmeans_points, "points_total": world * a.kmeans_points,
                                                  "dims": 128, "k": kk, "tf32_tflops_per_gpu": 2.0 * a.kmeans_points * kk * 128 / (bms * 1e-3) / 1e12,
                                                  "total_count_ok": int(kmb.total_counts.sum().item()) == world * a.kmeans_points}
                    del bigp, kmb
                    torch.cuda.empty_cache()
                except Exception as ex:  # noqa: BLE001
                    extra["kmeans_named_size"] = {"error": f"{type(ex).__name__}: {ex}"}
        except Exception as ex:  # noqa: BLE001
            extra["kmeans"] = {"error": f"{type(ex).__name__}: {ex}"}

    def ratio(x, y):
        return (x / y) if (x and y) else None

    vs = None
    if not a.skip_baseline:
        bt = base.get("terasort", {}).get("GB_per_s")
        vs = ratio(gbs, bt)
        base["ratios_ours_over_baseline"] = {
            "terasort_kernel_path": vs,
            "terasort_e2e": ratio((e2e or {}).get("value"), base.get("terasort_e2e", {}).get("GB_per_s")),
            "groupby_sparse_keys": ratio((groupby or {}).get("value"), base.get("groupby_sparse_keys", {}).get("rows_per_s")),
            "groupby_dense_ids": ratio(((groupby or {}).get("dense_ids") or {}).get("value"), base.get("groupby_dense_ids", {}).get("rows_per_s")),
            "broadcast_join": ratio((extra.get("broadcast_join") or {}).get("value"), base.get("broadcast_join", {}).get("rows_per_s")),
            "kmeans_iteration": ratio(base.get("kmeans", {}).get("ms_per_iteration"), ((extra.get("kmeans") or {}).get("value") or 0) * 1e3),
        }
    if rank == 0:
        line = {
            "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": vs,
            "vs_baseline_note": "the reference publishes no numbers (BASELINE.md); this is value / the PyTorch+NCCL implementation of the same sort timed in this run (BASELINE.json names it as the baseline)",
            "dtype": "exact (uint8 records, 80-bit keys); no reduced precision", "data": "synthetic",
            "impl": "ours", "valid": bool(ok),
            "config": {"model": "TeraSort-style RangePartition+OrderBy, 100-byte records, 10-byte keys",
                       "global_batch": world * n, "rows_per_gpu": n, "seq_len": W,
                       "parallelism": f"dp{world} (range-partitioned, one partition per GPU)",
                       "l2_policy": f"inputs exceed L2 ({n * W / 1e9:.1f} GB per GPU vs 126 MB)"},
            "clocks": clocks, "gpu_launches": launches, "e2e": e2e, "groupby_agg": groupby, "extra": extra,
            "baseline": None if a.skip_baseline else base,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
