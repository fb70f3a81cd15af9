"""Single-GPU program for ncu (round 2): one launch sequence of every hot kernel in its final form.
  fused shuffle (8 local partitions: the code path 8 ranks take, persistent TMA scatter), key stream
  scatter, radix sort (u32 pairs), key-only tie repair, row gather, hash aggregate on sparse 64-bit
  keys, k-means assign (tcgen05 CTA pairs) + certificate, fused join probe, tokenizer.
Usage: python bench/profile_r2.py [rows]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import joins, kmeans as KM, terasort, wordcount as WC  # noqa: E402
from dryad_b200.ops import core as K  # noqa: E402
from dryad_b200.parallel.dsort import RowSorter  # noqa: E402
from dryad_b200.parallel.kmeans import KMeans  # noqa: E402
from dryad_b200.parallel.shuffle import ColSpec, FusedShuffle  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
dev = torch.device("cuda", 0)
rows = terasort.generate_rows(n, 1)
seps = torch.tensor([(i + 1) * (2**64 // 8) - 2**63 for i in range(7)], dtype=torch.int64, device=dev) ^ (-2**63)
arena = SymmetricArena(int(n * 108 * 1.3) + (64 << 20))
sh = FusedShuffle(arena, [ColSpec(100, torch.uint8, 100)], int(n / 8 * 1.2) + 4096, parts_per_rank=8, emit_key=True, max_rows_in=n)
side = torch.cuda.Stream()
for _ in range(2):
    evs = sh.run_passes(rows, mode="range_asc", key_off=0, separators=seps, side=side, payload_ctas=148)
    torch.cuda.current_stream().wait_event(evs[-1])
sorter = RowSorter(SymmetricArena(1 << 20), n, 100, key_len=10)
for _ in range(2):
    out, m = sorter.sort(rows)
gk = K.hash_u64(torch.randint(0, 1 << 20, (n,), dtype=torch.int64, device=dev))
gv = torch.randint(0, 100, (n,), dtype=torch.int64, device=dev)
tbl = K.HashAggTable(K.agg_capacity(1 << 20), ["sum_i64", "count"], dev)
for _ in range(2):
    tbl.reset()
    tbl.update(gk, [gv, None])
pts, centres = KM.generate(2_000_000, 1024, 5, dev)
km = KMeans(SymmetricArena(8 << 20), 1024, 2_000_000)
km.set_centroids(centres + 0.05)
for _ in range(2):
    km.iterate(pts)
jr = joins.bench(0, 1, dev, n, 1 << 20, steps=1, warmup=1, max_over_ranks=lambda x: x, barrier=torch.cuda.synchronize)
text = torch.frombuffer(bytearray(WC.generate_text(200_000)), dtype=torch.uint8).to(dev)
for _ in range(2):
    WC.gpu_wordcount(text)
torch.cuda.synchronize()
print("profile target done", int(sh.status.item()), m, jr["valid"])
