"""Small-size run of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck).
Usage: compute-sanitizer --tool memcheck python bench/sanitize_gpu.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import kmeans as KM, terasort, wordcount as WC  # noqa: E402
from dryad_b200.ops import core as K  # noqa: E402
from dryad_b200.parallel.dsort import RowSorter  # noqa: E402
from dryad_b200.parallel.kmeans import KMeans  # noqa: E402
from dryad_b200.parallel.shuffle import ColSpec, FusedShuffle  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

dev = torch.device("cuda", 0)
n = 60_000
rows = terasort.generate_rows(n, 1)
out, m = RowSorter(SymmetricArena(1 << 20), n, 100, key_len=10).sort(rows)
assert int(K.check_sorted_rows(out, 0, 10).cpu()[0]) == 0
seps = torch.tensor([(i + 1) * (2**64 // 4) - 2**63 for i in range(3)], dtype=torch.int64, device=dev) ^ (-2**63)
arena = SymmetricArena(n * 108 * 2 + (8 << 20))
sh = FusedShuffle(arena, [ColSpec(100, torch.uint8, 100)], n // 2, parts_per_rank=4, emit_key=True, max_rows_in=n)
side = torch.cuda.Stream()
evs = sh.run_passes(rows, mode="range_asc", key_off=0, separators=seps, side=side, payload_ctas=16)
torch.cuda.current_stream().wait_event(evs[-1])
assert sum(sh.received_rows()) == n
keys = torch.randint(-2**62, 2**62, (n,), dtype=torch.int64, device=dev)
sk, sv = K.radix_sort_pairs(keys.clone(), None, iota=True)
gk = K.hash_u64(torch.randint(0, 5000, (n,), dtype=torch.int64, device=dev))
t = K.HashAggTable(K.agg_capacity(5000), ["sum_i64", "count", "min_i64"], dev)
t.update(gk, [torch.ones_like(gk), None, gk])
t.extract()
x, tot = K.exclusive_scan_i64(torch.randint(0, 9, (n,), dtype=torch.int64, device=dev))
text = torch.frombuffer(bytearray(WC.generate_text(2000)), dtype=torch.uint8).to(dev)
assert WC.gpu_wordcount(text)[4] == 0
pts, centres = KM.generate(20_000, 256, 5, dev)
km = KMeans(SymmetricArena(8 << 20), 256, 20_000)
km.set_centroids(centres + 0.05)
km.iterate(pts)
jt = K.JoinTable(torch.arange(1000, dtype=torch.int64, device=dev))
jt.probe(torch.randint(0, 1500, (n,), dtype=torch.int64, device=dev))
torch.cuda.synchronize()
print("sanitize target done")
