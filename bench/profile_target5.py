"""ncu target 5: the aggregate kernels — hashed (bucketised, lane-paired) and direct-addressed."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import groupby as GB  # noqa: E402
from dryad_b200.ops import core as K  # noqa: E402

dev = torch.device("cuda", 0)
n, nkeys = 40_000_000, 1 << 20
keys, vals = GB.generate(n, nkeys, 11, dev)
for _ in range(2):
    h = K.HashAggTable(K.agg_capacity(nkeys), ["sum_i64", "count"], dev)
    h.update(keys, [vals, None])
    gk, gl = h.extract()
    mm = K.minmax_i64(keys).tolist()
    d = K.DirectAggTable(mm[0], mm[1] - mm[0] + 1, ["sum_i64", "count"], dev)
    d.update(keys, [vals, None])
    dk, dl = d.extract()
torch.cuda.synchronize()
assert int(gl[1].sum()) == n and int(dl[1].sum()) == n and int(gl[0].sum()) == int(dl[0].sum())
print("ok", gk.numel(), dk.numel())
