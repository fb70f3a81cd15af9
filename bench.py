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

    # ---- end to end through the public API -------------------------------------------------------
    e2e = None
    if not a.skip_e2e:
        try:
            host_rows = torch.empty((n, W), dtype=torch.uint8).pin_memory()
            host_rows.copy_(rows)
            del sorter, arena, out
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
                   "h2d_bytes_per_step": n * W, "d2h_bytes_per_step": 24, "cpus_bound": len(bound) if bound else 0, "box_h2d_GBps_per_gpu": round(h2d_rate, 1),
                   "violations": int(summ[0]), "rows_out_local": int(summ[2])}
        except Exception as ex:  # noqa: BLE001
            e2e = {"error": f"{type(ex).__name__}: {ex}"}

    # ---- second headline metric: GroupBy-aggregate rows/s -----------------------------------------
    groupby = None
    if not a.skip_groupby:
        try:
            from dryad_b200.models import groupby as gbm
            groupby = gbm.bench(rank, world, dev, a.groupby_rows_per_gpu, a.groupby_keys, steps=a.steps,
                                warmup=a.warmup, max_over_ranks=max_over_ranks, barrier=barrier)
        except Exception as ex:  # noqa: BLE001
            groupby = {"error": f"{type(ex).__name__}: {ex}"}
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
        except Exception as ex:  # noqa: BLE001
            extra["broadcast_join"] = {"error": f"{type(ex).__name__}: {ex}"}
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
            extra["kmeans"] = {"metric": "kmeans_seconds_per_iteration", "value": kms * 1e-3, "unit": "s",
                               "points_per_gpu": npts, "dims": 128, "k": kk, "points_per_s": world * npts / (kms * 1e-3),
                               "tf32_tflops_per_gpu": 2.0 * npts * kk * 128 / (kms * 1e-3) / 1e12,
                               "total_count_ok": int(km.total_counts.sum().item()) == world * npts}
        except Exception as ex:  # noqa: BLE001
            extra["kmeans"] = {"error": f"{type(ex).__name__}: {ex}"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "exact (uint8 records, 80-bit keys); no reduced precision", "data": "synthetic",
            "impl": "ours", "valid": bool(ok),
            "config": {"model": "TeraSort-style RangePartition+OrderBy, 100-byte records, 10-byte keys",
                       "global_batch": world * n, "rows_per_gpu": n, "seq_len": W,
                       "parallelism": f"dp{world} (range-partitioned, one partition per GPU)",
                       "l2_policy": f"inputs exceed L2 ({n * W / 1e9:.1f} GB per GPU vs 126 MB)"},
            "clocks": clocks, "gpu_launches": launches, "e2e": e2e, "groupby_agg": groupby, "extra": extra,
        }
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
