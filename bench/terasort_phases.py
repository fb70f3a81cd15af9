"""Per-phase device times of the distributed TeraSort step (diagnostic; run under torchrun)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.models import terasort  # noqa: E402
from dryad_b200.parallel.dsort import RowSorter  # noqa: E402
from dryad_b200.parallel.symm import SymmetricArena  # noqa: E402

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 40_000_000
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
rows = terasort.generate_rows(n, 7, dev, rank)
arena = SymmetricArena((int(n * 1.12) + 8192) * 108 + (64 << 20) if world > 1 else 1 << 20, rank, world, None, dev)
s = RowSorter(arena, n, 100, key_len=10)
for _ in range(3):
    s.sort(rows)
s.profile = True
acc = {}
for _ in range(5):
    if world > 1:
        dist.barrier()
    s.sort(rows)
    for k, v in s.timing.items():
        acc[k] = acc.get(k, 0.0) + v / 5
t = torch.tensor(list(acc.values()), dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"n_gpus": world, "rows_per_gpu": n, "phase_ms_max_over_ranks": dict(zip(acc, [round(x, 3) for x in t.tolist()]))}))
if world > 1:
    dist.destroy_process_group()
