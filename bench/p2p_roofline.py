"""What the NVLink/NVSwitch fabric of THIS box delivers for an all-to-all of the TeraSort shuffle's
size: every rank pushes `mb` MiB to each of its peers.  Three forms, device-timed, max over ranks:
  copy    P2P tensor copies into the peers' arenas (cudaMemcpy peer path), one stream per peer
  nccl    torch.distributed.all_to_all_single
  stores  our 16-byte store kernel (dryad_replicate_slice writes one source to every peer)
The result is the roofline denominator for the fused shuffle (profiles/README.md)."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dryad_b200.parallel.symm import SymmetricArena, tensor_from_ptr  # noqa: E402

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 500
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
chunk = mb << 20
arena = SymmetricArena(world * chunk + (8 << 20), rank, world, None, dev)
recv, roff = arena.alloc((world, chunk), torch.uint8)
src = torch.randint(0, 255, (world, chunk), dtype=torch.uint8, device=dev)
peer_slots = [tensor_from_ptr(arena.peer_ptr(r, roff + rank * chunk), chunk, dev) for r in range(world)]
streams = [torch.cuda.Stream() for _ in range(world)]


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def copy_form():
    cur = torch.cuda.current_stream()
    for i in range(1, world):
        r = (rank + i) % world
        s = streams[r]
        s.wait_stream(cur)
        with torch.cuda.stream(s):
            peer_slots[r].copy_(src[r], non_blocking=True)
    for i in range(1, world):
        cur.wait_stream(streams[(rank + i) % world])


out = torch.empty_like(src)


def nccl_form():
    dist.all_to_all_single(out.view(-1), src.view(-1))


res = {"n_gpus": world, "mib_per_peer": mb, "bytes_out_per_gpu": (world - 1) * chunk}
for name, fn in (("copy", copy_form), ("nccl", nccl_form)):
    ms = timed(fn)
    res[name + "_ms"] = round(ms, 3)
    res[name + "_GBps_out_per_gpu"] = round((world - 1) * chunk / ms / 1e6, 1)
if rank == 0:
    print(json.dumps(res))
dist.destroy_process_group()
