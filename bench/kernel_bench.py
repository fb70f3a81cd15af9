"""Per-kernel device timings (CUDA events, warm-up, L2-exceeding inputs) with roofline fractions
against MEASURED_PEAKS.json.  Usage: python bench/kernel_bench.py [--rows N] [--json out]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import terasort  # noqa: E402
from dryad_b200.ops import core as K  # noqa: E402
from dryad_b200.parallel.dsort import RowSorter  # noqa: E402
from dryad_b200.parallel.shuffle import ColSpec, FusedShuffle  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402


def timeit(fn, iters=5, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=32_000_000)
    ap.add_argument("--json", default="")
    a = ap.parse_args()
    peaks = {"hbm_gbs": 6538.3}
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peaks = json.load(open(p))
    hbm = peaks["hbm_gbs"]
    n = a.rows
    res = {}
    dev = "cuda"

    keys = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device=dev)
    ka, kb = keys.clone(), torch.empty_like(keys)
    va, vb = torch.empty(n, dtype=torch.int32, device=dev), torch.empty(n, dtype=torch.int32, device=dev)

    def sort_pairs():
        ka.copy_(keys)
        K.radix_sort_pairs(ka, va, iota=True, keys_alt=kb, vals_alt=vb)
    t_copy = timeit(lambda: ka.copy_(keys))
    t = timeit(sort_pairs) - t_copy
    res["radix_sort_pairs_u64_ms"] = t
    res["radix_sort_pairs_u64_Gkeys_s"] = n / t / 1e6
    res["radix_sort_bytes_frac_of_measured_hbm"] = (n * 24 * 8 + n * 8) / (t * 1e-3) / 1e9 / hbm
    t = timeit(lambda: torch.sort(keys))
    res["torch_sort_u64_ms"] = t

    rows = terasort.generate_rows(n, 1)
    idx = torch.randperm(n, device=dev).to(torch.int32)
    out = torch.empty_like(rows)
    t = timeit(lambda: K.gather_rows(rows, idx, out))
    res["gather_rows_100B_ms"] = t
    res["gather_rows_GBs"] = n * 200 / t / 1e6
    res["gather_rows_frac_of_measured_hbm"] = n * 200 / (t * 1e-3) / 1e9 / hbm

    seps = torch.tensor([(i + 1) * (2**64 // 8) - 2**63 for i in range(7)], dtype=torch.int64, device=dev) ^ (-2**63)
    # 8 local partitions exercise the same code path as 8 ranks, writing to local HBM
    try:
        sh2 = FusedShuffle(SymmetricArena(int(n * 108 * 1.3 / 1) + (64 << 20)), [ColSpec(100, torch.uint8, 100)],
                           int(n / 8 * 1.2) + 4096, parts_per_rank=8, emit_key=True, max_rows_in=n)
        def shuffle2():
            sh2.run([rows], mode="range_asc", key_col=0, key_kind=1, key_off=0, separators=seps)
        t = timeit(shuffle2)
        st = int(sh2.status.item())
        res["shuffle_8parts_local_ms"] = t
        res["shuffle_status"] = st
        res["shuffle_GBs_payload"] = n * 100 / t / 1e6
        res["shuffle_frac_of_measured_hbm"] = (n * (100 + 100 + 8 + 10)) / (t * 1e-3) / 1e9 / hbm
    except Exception as ex:  # noqa: BLE001
        res["shuffle_error"] = repr(ex)

    arena1 = SymmetricArena(1 << 20)
    sorter = RowSorter(arena1, n, 100, key_len=10)
    t = timeit(lambda: sorter.sort(rows))
    res["terasort_1gpu_ms"] = t
    res["terasort_1gpu_GBs"] = n * 100 / t / 1e6
    o, m = sorter.sort(rows)
    chk = K.check_sorted_rows(o, 0, 10).cpu()
    res["terasort_violations"] = int(chk[0])

    card = 1 << 20
    gk = torch.randint(0, card, (n,), dtype=torch.int64, device=dev)
    gv = torch.randint(0, 100, (n,), dtype=torch.int64, device=dev)
    tbl = K.HashAggTable(card * 2, ["sum_i64", "count"], dev)
    def agg():
        tbl.reset()
        tbl.update(gk, [gv, None])
    t = timeit(agg)
    res["hash_agg_sum_count_ms"] = t
    res["hash_agg_Grows_s"] = n / t / 1e6
    res["hash_agg_frac_of_measured_hbm_read"] = n * 16 / (t * 1e-3) / 1e9 / hbm
    def torch_agg():
        torch.zeros(card, dtype=torch.int64, device=dev).index_add_(0, gk, gv)
        torch.bincount(gk, minlength=card)
    res["torch_index_add_bincount_ms"] = timeit(torch_agg)

    # k-means iteration: tcgen05 kernel vs PyTorch (cuBLAS TF32 matmul + argmin + index_add)
    try:
        from dryad_b200.models import kmeans as KM
        from dryad_b200.parallel.kmeans import KMeans
        del rows, out, idx, keys, ka, kb, va, vb, gk, gv
        torch.cuda.empty_cache()
        npts, kk = min(n, 8_000_000), 1024
        pts, centres = KM.generate(npts, kk, 5, dev)
        km = KMeans(SymmetricArena(8 << 20), kk, npts)
        km.set_centroids(centres + 0.05 * torch.randn_like(centres))
        t = timeit(lambda: km.iterate(pts), iters=3, warmup=2)
        res["kmeans_iter_ms"] = t
        res["kmeans_points"] = npts
        res["kmeans_Mpoints_s"] = npts / t / 1e3
        res["kmeans_tflops_tf32"] = 2.0 * npts * kk * 128 / (t * 1e-3) / 1e12
        res["kmeans_frac_of_measured_hbm"] = npts * 512 / (t * 1e-3) / 1e9 / hbm
        torch.backends.cuda.matmul.allow_tf32 = True
        c0 = km.centroids.clone()
        res["kmeans_torch_cublas_tf32_ms"] = timeit(lambda: KM.reference_iteration(pts, c0), iters=2, warmup=1)
    except Exception as ex:  # noqa: BLE001
        res["kmeans_error"] = repr(ex)

    for k, v in res.items():
        print(f"{k}: {v:.4f}" if isinstance(v, float) else f"{k}: {v}")
    if a.json:
        json.dump(res, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
