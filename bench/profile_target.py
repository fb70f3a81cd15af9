"""Short single-GPU program for ncu: one fused shuffle (8 local partitions, the same code path 8
ranks take), one radix sort, one gather, one hash aggregate.  Usage: python bench/profile_target.py [rows]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import terasort  # noqa: E402
from dryad_b200.ops import core as K  # noqa: E402
from dryad_b200.parallel.dsort import RowSorter  # noqa: E402
from dryad_b200.parallel.shuffle import ColSpec, FusedShuffle  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
dev = "cuda"
rows = terasort.generate_rows(n, 1)
seps = torch.tensor([(i + 1) * (2**64 // 8) - 2**63 for i in range(7)], dtype=torch.int64, device=dev) ^ (-2**63)
arena = SymmetricArena(int(n * 108 * 1.3) + (64 << 20))
sh = FusedShuffle(arena, [ColSpec(100, torch.uint8, 100)], int(n / 8 * 1.2) + 4096, parts_per_rank=8, emit_key=True, max_rows_in=n)
for _ in range(2):
    sh.run([rows], mode="range_asc", key_col=0, key_kind=1, key_off=0, separators=seps)
sorter = RowSorter(SymmetricArena(1 << 20), n, 100, key_len=10)
for _ in range(2):
    out, m = sorter.sort(rows)
gk = torch.randint(0, 1 << 20, (n,), dtype=torch.int64, device=dev)
gv = torch.randint(0, 100, (n,), dtype=torch.int64, device=dev)
tbl = K.HashAggTable(1 << 21, ["sum_i64", "count"], dev)
for _ in range(2):
    tbl.reset()
    tbl.update(gk, [gv, None])
torch.cuda.synchronize()
print("profile target done", int(sh.status.item()), m)
