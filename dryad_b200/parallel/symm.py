"""Symmetric memory: one peer-mapped HBM arena per rank (CUDA IPC over NVLink / NVSwitch).

The reference moves data between vertices as files behind an HTTP file server
(ClusterInterface/HttpScheduler.cs:65-91, ProcessService/HttpServer.cs:622-707).  On one NVSwitch
domain the equivalent "channel" is memory every GPU can address: each rank cudaMallocs one arena,
exports it with cudaIpcGetMemHandle, and opens every peer's handle.  Sub-allocation is a bump
allocator driven by identical call sequences on all ranks, so an allocation has the same offset
everywhere and `peer_ptr(rank, offset)` is pure arithmetic.  `torch.distributed` is used only to
exchange the 64-byte handles (control plane).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence, Tuple

import torch

from .. import _native as N
from ..ops import core as K

_TORCH_TYPESTR = {
    torch.uint8: "|u1", torch.int8: "|i1", torch.int16: "<i2", torch.int32: "<i4",
    torch.int64: "<i8", torch.float32: "<f4", torch.float64: "<f8", torch.bfloat16: None,
    torch.float16: "<f2", torch.bool: "|b1",
}


class _RawCuda:
    """Adapter exposing a raw device pointer through __cuda_array_interface__ (zero copy)."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None,
        }
        self._owner = owner


def tensor_from_ptr(ptr: int, nbytes: int, device, owner=None) -> torch.Tensor:
    if nbytes == 0:
        return torch.empty(0, dtype=torch.uint8, device=device)
    return torch.as_tensor(_RawCuda(ptr, nbytes, owner), device=device)


class SymmetricArena:
    """A symmetric HBM arena.  world == 1 needs no IPC and no process group."""

    def __init__(self, nbytes: int, rank: int = 0, world: int = 1, group=None, device=None):
        self.rank, self.world, self.group = rank, world, group
        d = torch.device(device if device is not None else "cuda")
        if d.index is None:
            d = torch.device("cuda", torch.cuda.current_device())
        self.device = d
        self.nbytes = int((nbytes + 511) // 512 * 512)
        lib = N.kernels()
        torch.cuda.set_device(self.device)
        base = C.c_void_p()
        N.check(lib.dryad_symm_alloc(self.nbytes, C.byref(base)), "symm_alloc")
        self.base = base.value
        self.peer_base: List[int] = [0] * world
        self.peer_base[rank] = self.base
        self._opened: List[int] = []
        if world > 1:
            import torch.distributed as dist
            h = C.create_string_buffer(64)
            N.check(lib.dryad_ipc_export(self.base, h), "ipc_export")
            handles: List[Optional[bytes]] = [None] * world
            dist.all_gather_object(handles, bytes(h.raw), group=group)
            for r in range(world):
                if r == rank:
                    continue
                p = C.c_void_p()
                N.check(lib.dryad_ipc_open(handles[r], C.byref(p)), f"ipc_open(rank {r})")
                self.peer_base[r] = p.value
                self._opened.append(p.value)
        self._cursor = 0
        self._whole = tensor_from_ptr(self.base, self.nbytes, self.device, owner=self)
        # control block: flags [FLAG_SLOTS][8] u64, then a status word
        self.flags_off = self.alloc_bytes(N.FLAG_SLOTS * N.MAX_PEERS * 8)
        self.epoch = 0

    # -- allocation ---------------------------------------------------------------------------
    def alloc_bytes(self, nbytes: int, align: int = 512) -> int:
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise MemoryError(f"symmetric arena exhausted: need {off + nbytes}, have {self.nbytes}")
        self._cursor = off + nbytes
        return off

    def check_symmetric(self, what: str = "") -> None:
        """Init-time guard: every rank must have made the same allocations (same cursor), or peer
        addresses computed from local offsets would point at the wrong objects."""
        if self.world == 1:
            return
        import torch.distributed as dist
        cursors = [None] * self.world
        dist.all_gather_object(cursors, self._cursor, group=self.group)
        if len(set(cursors)) != 1:
            raise RuntimeError(f"symmetric arena diverged across ranks {cursors} ({what}): "
                               "allocate with identical sizes on every rank")

    def mark(self) -> int:
        return self._cursor

    def release_to(self, mark: int) -> None:
        """Stack discipline: drop every allocation made after `mark` (same call on all ranks)."""
        self._cursor = mark

    def view(self, off: int, shape: Sequence[int], dtype: torch.dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= int(s)
        nb = n * torch.empty(0, dtype=dtype).element_size()
        return self._whole[off:off + nb].view(dtype).view(*shape)

    def alloc(self, shape: Sequence[int], dtype: torch.dtype) -> Tuple[torch.Tensor, int]:
        n = 1
        for s in shape:
            n *= int(s)
        off = self.alloc_bytes(max(n, 1) * torch.empty(0, dtype=dtype).element_size())
        return self.view(off, shape, dtype), off

    def peer_ptr(self, r: int, off: int) -> int:
        return self.peer_base[r] + off

    def peer_ptr_table(self, off: int) -> torch.Tensor:
        """Device int64[world] with every rank's address of the allocation at `off`."""
        return torch.tensor([self.peer_ptr(r, off) for r in range(self.world)], dtype=torch.int64).to(self.device)

    # -- synchronisation -----------------------------------------------------------------------
    def next_epoch(self) -> int:
        self.epoch += 1
        return self.epoch

    def flags_local_ptr(self) -> int:
        return self.base + self.flags_off

    def flags_peers_table(self) -> torch.Tensor:
        if not hasattr(self, "_flags_tbl"):
            self._flags_tbl = self.peer_ptr_table(self.flags_off)
        return self._flags_tbl

    def barrier(self, slot: int = 2, status: Optional[torch.Tensor] = None) -> None:
        """Device-side epoch barrier on the current stream (no host sync, no NCCL)."""
        if status is None:
            if not hasattr(self, "_status"):
                self._status = torch.zeros(1, dtype=torch.int32, device=self.device)
            status = self._status
        e = self.next_epoch()
        N.check(N.kernels().dryad_signal_barrier(K.ptr(self.flags_peers_table()), self.flags_local_ptr(),
                                                 self.rank, self.world, slot, e, K.ptr(status),
                                                 K.stream()), "signal_barrier")
        K._count()

    def close(self) -> None:
        lib = N.kernels()
        torch.cuda.synchronize(self.device)
        for p in self._opened:
            lib.dryad_ipc_close(p)
        self._opened = []
        if self.base:
            self._whole = None
            lib.dryad_symm_free(self.base)
            self.base = 0
