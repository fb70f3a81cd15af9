"""NVLS buffers: symmetric memory with an NVSwitch MULTICAST mapping.

A store to the multicast address lands in every GPU's copy of the buffer; a `multimem.ld_reduce`
returns the sum of all copies, computed inside the switch (csrc/kernels/nvls.cu).  Creating the
multicast object (cuMulticastCreate / bind / map across processes) is control-plane work that
torch's symmetric-memory rendezvous already does; this module only wraps it and hands raw pointers
to our kernels.  `NvlsBuffer.create` returns None when the platform has no multicast support — the
callers then keep their P2P forms (peer loads / replicated stores).
"""
from __future__ import annotations

import os
from typing import List, Optional

import torch

from .. import _native as N
from ..ops import core as K


class NvlsBuffer:
    def __init__(self, tensor: torch.Tensor, handle, rank: int, world: int):
        self.t, self.h, self.rank, self.world = tensor, handle, rank, world
        self.nbytes = tensor.numel() * tensor.element_size()
        self.mc_ptr: int = int(handle.multicast_ptr)
        self.peer_ptrs: List[int] = [int(p) for p in handle.buffer_ptrs]
        self._cursor = 0

    @staticmethod
    def create(nbytes: int, device, rank: int, world: int) -> Optional["NvlsBuffer"]:
        if world < 2 or os.environ.get("DRYAD_NVLS", "1") == "0":
            return None
        try:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm
            if not dist.is_initialized():
                return None
            t = symm.empty(int((nbytes + 4095) // 4096 * 4096), dtype=torch.uint8, device=device)
            h = symm.rendezvous(t, dist.group.WORLD.group_name)
            if not int(getattr(h, "multicast_ptr", 0)):
                return None
            return NvlsBuffer(t, h, rank, world)
        except Exception:  # noqa: BLE001  (no fabric / driver support: P2P forms stay in use)
            return None

    # identical call sequences on all ranks give identical offsets, as in SymmetricArena
    def alloc(self, shape, dtype: torch.dtype, align: int = 256):
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.nbytes:
            raise MemoryError("NVLS buffer exhausted")
        self._cursor = off + nbytes
        view = self.t[off:off + nbytes].view(dtype).view(*shape)
        return view, off

    def mc(self, off: int) -> int:
        return self.mc_ptr + off

    # ---- data plane ----------------------------------------------------------------------------
    def allreduce_f32(self, off: int, n: int) -> None:
        N.check(N.kernels().dryad_nvls_allreduce_f32(self.mc(off), n, self.rank, self.world, K.stream()), "nvls_allreduce_f32")
        K._count()

    def allreduce_u64(self, off: int, n: int) -> None:
        N.check(N.kernels().dryad_nvls_allreduce_u64(self.mc(off), n, self.rank, self.world, K.stream()), "nvls_allreduce_u64")
        K._count()

    def broadcast(self, dst_off: int, src: torch.Tensor) -> None:
        """Multicast `src` (local, contiguous, 16-byte multiple) into byte offset dst_off of every copy."""
        K.require_cuda(src)
        N.check(N.kernels().dryad_nvls_broadcast(self.mc(dst_off), K.ptr(src), src.numel() * src.element_size(), K.stream()),
                "nvls_broadcast")
        K._count()
