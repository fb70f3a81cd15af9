"""Per-phase device times of the distributed GroupBy Sum/Count step (diagnostic; run alone or under
torchrun)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import groupby  # noqa: E402
from dryad_b200.ops import core as K  # noqa: E402
from dryad_b200.parallel.dagg import DistributedHashAggregate  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000_000
nkeys = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
keys, vals = groupby.generate(n, nkeys, 11, dev, rank, sparse=(len(sys.argv) > 3 and sys.argv[3] == "sparse"))
fcap = K.agg_capacity(int(1.25 * nkeys) // world + 1024)
arena = SymmetricArena(fcap * 8 * 3 + (fcap // 3 + 1024) * world * 24 + (16 << 20), rank, world, None, dev)
agg = DistributedHashAggregate(arena, ["sum_i64", "count"], K.agg_capacity(nkeys), fcap)
for _ in range(3):
    agg.run(keys, [vals, None])


def ev():
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


acc = {}
R = 5
for _ in range(R):
    if world > 1:
        dist.barrier()
    marks = [("start", ev())]
    agg.final.reset()
    if world > 1:
        agg.local.reset()
    marks.append(("reset", ev()))
    (agg.local if world > 1 else agg.final).update(keys, [vals, None])
    marks.append(("local update", ev()))
    if world > 1 and agg.shuffle is not None:
        agg.nrows.zero_()
        agg.local.extract_rows(agg.rows, agg.nrows)
        m = int(agg.nrows.item())
        marks.append(("extract partial rows", ev()))
        agg.shuffle.run([agg.rows[:m]], mode="hash", key_col=0, key_kind=0)
        got = agg.shuffle.received_rows()[0]
        marks.append(("shuffle partials", ev()))
        agg.final.update_rows(agg.shuffle.recv[0][0], got, combine=True)
        marks.append(("fold received", ev()))
    elif world > 1:
        arena.barrier(slot=2)
        marks.append(("barrier", ev()))
        agg.local.merge_into_peers(world, agg.peer_keys, agg.peer_vals, agg.final_cap)
        marks.append(("merge into owners", ev()))
        arena.barrier(slot=3)
        marks.append(("barrier 2", ev()))
    gk, gl = agg.final.extract()
    marks.append(("extract", ev()))
    torch.cuda.synchronize()
    for (_, a), (nm, b) in zip(marks, marks[1:]):
        acc[nm] = acc.get(nm, 0.0) + a.elapsed_time(b) / R
t = torch.tensor(list(acc.values()), dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"n_gpus": world, "rows_per_gpu": n, "distinct_keys": nkeys, "groups_out_rank0": int(gk.numel()),
                      "phase_ms_max_over_ranks": dict(zip(acc, [round(x, 3) for x in t.tolist()]))}))
if world > 1:
    dist.destroy_process_group()
