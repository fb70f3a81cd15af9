"""GroupBy Sum+Count throughput vs number of distinct keys (front table on/off)."""
import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.ops import core as K
dev = torch.device("cuda")
n = 100_000_000
g = torch.Generator(device=dev).manual_seed(1)
vals = torch.randint(0, 1000, (n,), dtype=torch.int64, device=dev, generator=g)
out = {}
for nk in (4, 64, 1024, 4096, 65536, 1 << 20):
    keys = torch.randint(0, nk, (n,), dtype=torch.int64, device=dev, generator=g) * 0x9E3779B97F4A7C15 % (1 << 62)
    t = K.HashAggTable(K.agg_capacity(nk), ["sum_i64", "count"], dev)
    for _ in range(2):
        t.reset(); t.update(keys, [vals, None])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        t.reset(); t.update(keys, [vals, None])
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    gk, gl = t.extract()
    assert int(gl[1].sum()) == n
    out[str(nk)] = round(n / ms / 1e6, 1)
print(json.dumps({"front": os.environ.get("DRYAD_AGG_FRONT", "1"), "Grows_per_s_by_distinct_keys": out}))
