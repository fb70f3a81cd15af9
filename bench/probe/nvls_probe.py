"""Does this box give us NVLS multicast memory through torch's symmetric-memory rendezvous?"""
import os
import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
dist.init_process_group("nccl", device_id=dev)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1 << 20, dtype=torch.uint8, device=dev)
    hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
    print(rank, "buffer_ptrs", [hex(p) for p in hdl.buffer_ptrs][:3], "multicast_ptr", hex(hdl.multicast_ptr),
          "signal_pad", hex(hdl.signal_pad_ptrs[0]) if hasattr(hdl, "signal_pad_ptrs") else None, flush=True)
    if hdl.multicast_ptr:
        # functional check with torch's own multimem op if present
        x = symm.empty(1024, dtype=torch.float32, device=dev)
        x.fill_(rank + 1)
        symm.rendezvous(x, dist.group.WORLD.group_name)
        dist.barrier()
        y = torch.ops.symm_mem.multimem_all_reduce_(x, "sum", dist.group.WORLD.group_name)
        torch.cuda.synchronize()
        print(rank, "multimem_all_reduce ->", float(y[0]), flush=True)
except Exception as ex:  # noqa: BLE001
    print(rank, "symmetric memory probe failed:", type(ex).__name__, str(ex)[:300], flush=True)
dist.barrier()
dist.destroy_process_group()
