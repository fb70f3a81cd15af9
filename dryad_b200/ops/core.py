"""Thin torch-tensor wrappers over the C ABI in include/dryad_b200.h.

Conventions: 64-bit keys travel as `torch.int64` tensors whose BIT PATTERN is the unsigned key
(torch has no general uint64 arithmetic); every wrapper launches on the current CUDA stream and
never synchronises unless it must return a host value.  A kernel launch counter is kept so
benchmarks can report how many of our kernels ran (`launch_count()`).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from .. import _native as N

_launches = 0
_scratch_cache: dict = {}

AGG = {"count": 0, "sum_i64": 1, "sum_i64_checked": 2, "sum_f64": 3, "min_i64": 4, "max_i64": 5,
       "min_f64": 6, "max_f64": 7}
RED = {"sum": 0, "min": 1, "max": 2}
DT = {torch.int64: 0, torch.float64: 1, torch.int32: 2, torch.float32: 3, torch.uint8: 4,
      torch.int16: 5, torch.bool: 4}
PART = {"hash": 0, "range_asc": 1, "range_desc": 2, "direct": 3}


def launch_count() -> int:
    return _launches


def _count(k: int = 1) -> None:
    global _launches
    _launches += k


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def require_cuda(*ts: torch.Tensor) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("dryad_b200.ops kernels need CUDA tensors (use the oracle on CPU)")
        if t is not None and not t.is_contiguous():
            raise ValueError("dryad_b200.ops kernels need contiguous tensors")


def scratch(nbytes: int, device, tag: str = "") -> torch.Tensor:
    """A cached scratch buffer of at least nbytes (one per (device, tag); grows monotonically).
    Callers on one stream may share it because kernels on a stream serialise."""
    key = (str(device), tag)
    buf = _scratch_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
        _scratch_cache[key] = buf
    return buf


# ---- K14 -------------------------------------------------------------------------------------
def hash_u64(keys: torch.Tensor) -> torch.Tensor:
    require_cuda(keys)
    out = torch.empty_like(keys)
    N.check(N.kernels().dryad_hash_u64(ptr(keys), ptr(out), keys.numel(), stream()), "hash_u64")
    _count()
    return out


def hash_bytes(data: torch.Tensor, offsets: torch.Tensor) -> torch.Tensor:
    require_cuda(data, offsets)
    n = offsets.numel() - 1
    out = torch.empty(n, dtype=torch.int64, device=data.device)
    N.check(N.kernels().dryad_hash_bytes(ptr(data), ptr(offsets), ptr(out), n, stream()), "hash_bytes")
    _count()
    return out


# ---- K4 --------------------------------------------------------------------------------------
def radix_sort_pairs(keys: torch.Tensor, vals: Optional[torch.Tensor] = None, *,
                     begin_bit: int = 0, end_bit: Optional[int] = None, iota: bool = False,
                     keys_alt: Optional[torch.Tensor] = None,
                     vals_alt: Optional[torch.Tensor] = None):
    """Stable ascending sort by the unsigned bit pattern of `keys` (int64 or int32 storage).
    Returns (sorted_keys, sorted_vals).  INPUT BUFFERS ARE CLOBBERED (ping-pong).  With
    iota=True the values start as 0..n-1 (a sort permutation) and `vals` may be None."""
    require_cuda(keys, vals)
    n = keys.numel()
    bits = keys.element_size() * 8
    end_bit = bits if end_bit is None else end_bit
    if keys.dtype not in (torch.int64, torch.int32):
        raise TypeError("radix_sort_pairs: keys must be int64 or int32 bit patterns")
    if iota and vals is None:
        vals = torch.empty(n, dtype=torch.int32, device=keys.device)
    if vals is not None and vals.dtype != torch.int32:
        raise TypeError("radix_sort_pairs: values must be int32 (u32 payload)")
    if keys_alt is None:
        keys_alt = torch.empty_like(keys)
    if vals is not None and vals_alt is None:
        vals_alt = torch.empty_like(vals)
    lib = N.kernels()
    sb = lib.dryad_radix_sort_scratch_bytes(n, begin_bit, end_bit)
    sc = scratch(sb, keys.device, "radix")
    fn = lib.dryad_radix_sort_pairs_u64 if bits == 64 else lib.dryad_radix_sort_pairs_u32
    in_alt = C.c_int(0)
    N.check(fn(ptr(keys), ptr(keys_alt), ptr(vals), ptr(vals_alt), n, begin_bit, end_bit,
               1 if iota else 0, ptr(sc), sc.numel(), C.byref(in_alt), stream()), "radix_sort")
    npass = (end_bit - begin_bit + 7) // 8
    _count(2 + npass if n else 0)
    if in_alt.value:
        return keys_alt, vals_alt
    return keys, vals


def sort_key_bits(x: torch.Tensor, descending: bool = False) -> torch.Tensor:
    """Order-preserving map of a numeric column onto unsigned 64-bit patterns (int64 storage)."""
    if x.dtype in (torch.int64, torch.int32, torch.int16, torch.int8):
        b = x.to(torch.int64) ^ torch.tensor(-0x8000000000000000, dtype=torch.int64, device=x.device)
    elif x.dtype in (torch.uint8, torch.bool):
        b = x.to(torch.int64)
    elif x.dtype in (torch.float64, torch.float32):
        v = x.to(torch.float64).view(torch.int64)
        neg = v < 0
        b = torch.where(neg, ~v, v ^ torch.tensor(-0x8000000000000000, dtype=torch.int64, device=x.device))
    else:
        raise TypeError(f"unsupported sort key dtype {x.dtype}")
    return ~b if descending else b


# ---- row helpers ------------------------------------------------------------------------------
def extract_be64(rows: torch.Tensor, key_off: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    require_cuda(rows)
    n, w = rows.shape
    if out is None:
        out = torch.empty(n, dtype=torch.int64, device=rows.device)
    N.check(N.kernels().dryad_extract_be64(ptr(rows), n, w, key_off, ptr(out), stream()), "extract_be64")
    _count()
    return out


def gather_rows(src: torch.Tensor, idx: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[i] = src[idx[i]] for a 1-D (4/8-byte) column or a 2-D uint8 row table."""
    require_cuda(src, idx)
    n = idx.numel()
    if src.dim() == 1:
        width = src.element_size()
        shape = (n,)
    else:
        width = src.shape[1] * src.element_size()
        shape = (n, src.shape[1])
    if width % 4:
        raise ValueError("gather_rows: row width must be a multiple of 4 bytes")
    if out is None:
        out = torch.empty(shape, dtype=src.dtype, device=src.device)
    N.check(N.kernels().dryad_gather_rows(ptr(src), ptr(idx), ptr(out), n, width, stream()), "gather_rows")
    _count()
    return out


def fix_ties(keys: torch.Tensor, idx: torch.Tensor, rows: torch.Tensor, extra_off: int, extra: int,
             *, max_run: int = 4096, descending: bool = False, prefix_mask: int = (1 << 64) - 1) -> torch.Tensor:
    """Order runs of equal 64-bit prefixes by the next `extra` key bytes.  Returns a device int32
    counter of runs longer than max_run (left untouched)."""
    require_cuda(keys, idx, rows)
    long_runs = torch.zeros(1, dtype=torch.int32, device=keys.device)
    N.check(N.kernels().dryad_fix_ties(ptr(keys), ptr(idx), keys.numel(), ptr(rows), rows.shape[1],
                                       extra_off, extra, max_run, int(descending), prefix_mask,
                                       ptr(long_runs), stream()), "fix_ties")
    _count()
    return long_runs


def check_sorted_rows(rows: torch.Tensor, key_off: int, key_len: int, descending: bool = False) -> torch.Tensor:
    """Device tensor [violations, wrapping sum of 64-bit key prefixes] (int64 storage)."""
    require_cuda(rows)
    res = torch.zeros(2, dtype=torch.int64, device=rows.device)
    N.check(N.kernels().dryad_check_sorted_rows(ptr(rows), rows.shape[0], rows.shape[1], key_off,
                                                key_len, int(descending), ptr(res), stream()),
            "check_sorted_rows")
    _count()
    return res


# ---- K3 --------------------------------------------------------------------------------------
def sample_keys(keys: torch.Tensor, seed: int, rate: float, cap: int):
    require_cuda(keys)
    out = torch.empty(cap, dtype=torch.int64, device=keys.device)
    cnt = torch.zeros(1, dtype=torch.int64, device=keys.device)
    N.check(N.kernels().dryad_sample_keys(ptr(keys), keys.numel(), seed, rate, ptr(out), ptr(cnt),
                                          cap, stream()), "sample_keys")
    _count()
    return out, cnt


def sample_rows(rows: torch.Tensor, key_off: int, seed: int, rate: float, cap: int):
    require_cuda(rows)
    out = torch.empty(cap, dtype=torch.int64, device=rows.device)
    cnt = torch.zeros(1, dtype=torch.int64, device=rows.device)
    N.check(N.kernels().dryad_sample_rows(ptr(rows), rows.shape[1], key_off, rows.shape[0], seed,
                                          rate, ptr(out), ptr(cnt), cap, stream()), "sample_rows")
    _count()
    return out, cnt


def select_separators(sorted_samples: torch.Tensor, count: int, pcount: int, descending: bool = False) -> torch.Tensor:
    require_cuda(sorted_samples)
    seps = torch.empty(max(pcount - 1, 0), dtype=torch.int64, device=sorted_samples.device)
    N.check(N.kernels().dryad_select_separators(ptr(sorted_samples), count, pcount, int(descending),
                                                ptr(seps), stream()), "select_separators")
    _count()
    return seps


# ---- K6 --------------------------------------------------------------------------------------
def _next_pow2(x: int) -> int:
    p = 1
    while p < x:
        p <<= 1
    return p


def agg_lane_stride(nlanes: int) -> int:
    """Words per slot of a HashAggTable value array (1, 2, 4 or 8)."""
    return 1 if nlanes <= 1 else 2 if nlanes <= 2 else 4 if nlanes <= 4 else 8


def agg_capacity(groups: int, load: float = 1.0 / 3.0) -> int:
    """Table capacity (a multiple of the 4-key bucket) for an expected number of groups.  One
    third full keeps second-bucket probes under 1 % (bench/micro/agg_variants3.cu)."""
    return max(64, (int(groups / load) + 3) // 4 * 4)


class HashAggTable:
    """Open-addressing aggregation table in HBM: int64 key bits -> up to 8 aggregate lanes.
    `keys[cap]` is probed in 4-key buckets; `vals[cap, L]` is slot-major (L = lane stride) so a
    group's aggregates share one 32-byte sector (see csrc/kernels/hash_table.cu)."""

    def __init__(self, capacity: int, ops: Sequence[str], device, *, keys=None, vals=None):
        self.cap = max((int(capacity) + 3) // 4 * 4, 64)
        self.ops = list(ops)
        self.nl = len(self.ops)
        self.stride = agg_lane_stride(self.nl)
        self.device = device
        self.keys = keys if keys is not None else torch.empty(self.cap, dtype=torch.int64, device=device)
        self.vals = vals if vals is not None else torch.empty((self.cap, self.stride), dtype=torch.int64, device=device)
        if self.keys.numel() != self.cap or self.vals.numel() != self.cap * self.stride:
            raise ValueError("HashAggTable: keys/vals do not match the capacity and lane stride")
        self._ops_c = (C.c_int * max(self.nl, 1))(*[AGG[o] for o in self.ops])
        self.status = torch.zeros(1, dtype=torch.int32, device=device)
        self._work = None
        self.reset()

    def reset(self):
        N.check(N.kernels().dryad_hash_agg_init(ptr(self.keys), self.cap, self.nl, self._ops_c,
                                                ptr(self.vals), stream()), "hash_agg_init")
        self.status.zero_()
        _count()

    def update(self, keys: torch.Tensor, cols: Sequence[Optional[torch.Tensor]], combine: bool = False):
        require_cuda(keys, *[c for c in cols if c is not None])
        ins = []
        for c in cols:
            if c is None:
                ins.append(None)
            else:
                if c.dtype not in (torch.int64, torch.float64):
                    raise TypeError("aggregate input lanes must be int64 or float64")
                ins.append(c.data_ptr())
        in_c = (C.c_void_p * max(self.nl, 1))(*ins) if self.nl else (C.c_void_p * 1)()
        if self._work is None:
            self._work = torch.zeros(4, dtype=torch.int32, device=self.device)
        N.check(N.kernels().dryad_hash_agg_update_auto(ptr(self.keys), self.cap, self.nl, self._ops_c,
                                                       ptr(self.vals), ptr(keys), in_c, keys.numel(),
                                                       int(combine), ptr(self.status), ptr(self._work), stream()),
                "hash_agg_update")
        _count(3 if keys.numel() >= (1 << 18) else 1)

    def merge_into_peers(self, world: int, peer_keys_ptrs: Sequence[int], peer_vals_ptrs: Sequence[int], peer_cap: int):
        pk = (C.c_void_p * world)(*peer_keys_ptrs)
        pv = (C.c_void_p * world)(*peer_vals_ptrs)
        N.check(N.kernels().dryad_hash_agg_merge_peers(ptr(self.keys), self.cap, self.nl, self._ops_c,
                                                       ptr(self.vals), world, pk, pv, peer_cap,
                                                       ptr(self.status), stream()),
                "hash_agg_merge_peers")
        _count()

    def update_rows(self, rows: torch.Tensor, n: int, combine: bool = True):
        """Fold packed rows [n, 1 + nlanes] int64 {key, values...} into the table."""
        require_cuda(rows)
        N.check(N.kernels().dryad_hash_agg_update_rows(ptr(self.keys), self.cap, self.nl, self._ops_c, ptr(self.vals),
                                                       ptr(rows), int(n), int(combine), ptr(self.status), stream()),
                "hash_agg_update_rows")
        _count()

    def extract_rows(self, out: torch.Tensor, count: torch.Tensor) -> None:
        """Compact the occupied slots into packed rows `out` [cap_rows, 1 + nlanes] int64; `count`
        (int64[1], zeroed by the caller) receives the number of rows.  No host sync."""
        N.check(N.kernels().dryad_hash_agg_extract_rows(ptr(self.keys), self.cap, self.nl, self._ops_c, ptr(self.vals),
                                                        ptr(out), ptr(count), out.shape[0], stream()),
                "hash_agg_extract_rows")
        _count()

    def extract(self, max_out: Optional[int] = None):
        """Returns (keys[m], lanes[nl][m]) of occupied slots (order unspecified).  Syncs once to
        learn m."""
        max_out = self.cap if max_out is None else max_out
        out_keys = torch.empty(max_out, dtype=torch.int64, device=self.device)
        out_lanes = torch.empty((max(self.nl, 1), max_out), dtype=torch.int64, device=self.device)
        lane_ptrs = torch.tensor([out_lanes[l].data_ptr() for l in range(max(self.nl, 1))], dtype=torch.int64).to(self.device)
        cnt = torch.zeros(1, dtype=torch.int64, device=self.device)
        N.check(N.kernels().dryad_hash_agg_extract(ptr(self.keys), self.cap, self.nl, self._ops_c,
                                                   ptr(self.vals), ptr(out_keys), ptr(lane_ptrs),
                                                   ptr(cnt), max_out, stream()), "hash_agg_extract")
        _count()
        m = int(cnt.item())
        st = int(self.status.item())
        if st == 1:
            raise N.NativeError("hash aggregate table full")
        if st == 3:
            raise OverflowError("checked int64 Sum overflowed")
        if m > max_out:
            raise N.NativeError("hash_agg_extract: output capacity too small")
        return out_keys[:m], out_lanes[:, :m]


def minmax_i64(col: torch.Tensor) -> torch.Tensor:
    """[min, max] of an int64 column in one pass (device tensor of 2; no host sync)."""
    require_cuda(col)
    out = torch.empty(2, dtype=torch.int64, device=col.device)
    N.check(N.kernels().dryad_minmax_i64(ptr(col), col.numel(), ptr(out), stream()), "minmax_i64")
    _count(2)
    return out


class DirectAggTable:
    """Aggregation table for a DENSE key range [key_lo, key_lo + cap): the key is the slot, so an
    update is one lane-paired reduction per row with no probe (csrc/kernels/hash_table.cu, direct
    mode).  Same extract / merge contract as HashAggTable; raw-row updates only."""

    def __init__(self, key_lo: int, span: int, ops: Sequence[str], device):
        self.key_lo = int(key_lo)
        self.cap = max((int(span) + 3) // 4 * 4, 64)
        self.ops = list(ops)
        self.nl = len(self.ops)
        self._ops_c = (C.c_int * max(self.nl, 1))(*[AGG[o] for o in self.ops])
        self.stride = N.kernels().dryad_hash_agg_direct_lane_stride(self.nl, self._ops_c)
        if self.stride < 1:
            raise ValueError("too many aggregate lanes for a direct table")
        self.device = device
        self.vals = torch.empty((self.cap, self.stride), dtype=torch.int64, device=device)
        self.status = torch.zeros(1, dtype=torch.int32, device=device)
        self._work = None
        self.reset()

    def _lo(self) -> int:
        return self.key_lo & ((1 << 64) - 1)

    def reset(self, key_lo: Optional[int] = None):
        if key_lo is not None:
            self.key_lo = int(key_lo)
        N.check(N.kernels().dryad_hash_agg_init_direct(ptr(self.vals), self.cap, self.nl, self._ops_c, stream()),
                "hash_agg_init_direct")
        self.status.zero_()
        _count()

    def update(self, keys: torch.Tensor, cols: Sequence[Optional[torch.Tensor]]):
        require_cuda(keys, *[c for c in cols if c is not None])
        in_c = (C.c_void_p * max(self.nl, 1))(*[None if c is None else c.data_ptr() for c in cols]) if self.nl else (C.c_void_p * 1)()
        if self._work is None:
            self._work = torch.zeros(4, dtype=torch.int32, device=self.device)
        N.check(N.kernels().dryad_hash_agg_update_direct_auto(ptr(self.vals), self.cap, self.nl, self._ops_c, self._lo(),
                                                              ptr(keys), in_c, keys.numel(), ptr(self.status),
                                                              ptr(self._work), stream()),
                "hash_agg_update_direct")
        _count(3 if keys.numel() >= (1 << 18) else 1)

    def merge_into_peers(self, world: int, peer_keys_ptrs: Sequence[int], peer_vals_ptrs: Sequence[int], peer_cap: int):
        pk = (C.c_void_p * world)(*peer_keys_ptrs)
        pv = (C.c_void_p * world)(*peer_vals_ptrs)
        N.check(N.kernels().dryad_hash_agg_merge_peers_direct(ptr(self.vals), self.cap, self.nl, self._ops_c, self._lo(),
                                                              world, pk, pv, peer_cap, ptr(self.status), stream()),
                "hash_agg_merge_peers_direct")
        _count()

    def extract(self, max_out: Optional[int] = None):
        max_out = self.cap if max_out is None else max_out
        out_keys = torch.empty(max_out, dtype=torch.int64, device=self.device)
        out_lanes = torch.empty((max(self.nl, 1), max_out), dtype=torch.int64, device=self.device)
        lane_ptrs = torch.tensor([out_lanes[l].data_ptr() for l in range(max(self.nl, 1))], dtype=torch.int64).to(self.device)
        cnt = torch.zeros(1, dtype=torch.int64, device=self.device)
        N.check(N.kernels().dryad_hash_agg_extract_direct(ptr(self.vals), self.cap, self.nl, self._ops_c, self._lo(),
                                                          ptr(out_keys), ptr(lane_ptrs), ptr(cnt), max_out, stream()),
                "hash_agg_extract_direct")
        _count()
        m = int(cnt.item())
        st = int(self.status.item())
        if st == 4:
            raise N.NativeError("direct aggregate table: key outside the declared range")
        if st == 3:
            raise OverflowError("checked int64 Sum overflowed")
        return out_keys[:m], out_lanes[:, :m]


# ---- K8 / K10 --------------------------------------------------------------------------------
class JoinTable:
    def __init__(self, build_keys: torch.Tensor, load: float = 0.5):
        require_cuda(build_keys)
        n = build_keys.numel()
        self.cap = _next_pow2(max(int(n / load) + 1, 64))
        dev = build_keys.device
        self.keys = torch.empty(self.cap, dtype=torch.int64, device=dev)
        self.head = torch.empty(self.cap, dtype=torch.int32, device=dev)
        self.next = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)
        self.n = n
        self._unique = None
        N.check(N.kernels().dryad_join_build(ptr(self.keys), ptr(self.head), self.cap, ptr(build_keys),
                                             ptr(self.next), n, ptr(self.status), stream()), "join_build")
        _count(2)

    @property
    def unique(self) -> bool:
        """True when every build key is distinct (syncs once): the build side is a dimension table
        and the probe can be fused into the consuming pipeline (`expr.probe`)."""
        if self._unique is None:
            st = int(self.status.item())
            if st == 1:
                raise N.NativeError("join table full")
            self._unique = st == 0
        return self._unique

    def slot_payload(self, col: torch.Tensor) -> torch.Tensor:
        """`col` (one value per build row) permuted into table-slot order: out[s] = col[head[s]]
        (unique-key tables; pairs with `expr.probe_slot`)."""
        return gather_rows(col.contiguous(), self.head.clamp(min=0).contiguous())

    def probe(self, probe_keys: torch.Tensor):
        """Returns (probe_idx, build_idx) int32 index pairs of all matches."""
        require_cuda(probe_keys)
        n = probe_keys.numel()
        dev = probe_keys.device
        counts = torch.empty(n, dtype=torch.int64, device=dev)
        lib = N.kernels()
        N.check(lib.dryad_join_probe(ptr(self.keys), ptr(self.head), self.cap, ptr(self.next),
                                     ptr(probe_keys), n, 0, ptr(counts), None, None, stream()), "join_probe")
        _count()
        offs, tot = exclusive_scan_i64(counts)                 # counts -> output offsets, in place
        total = int(tot.item()) if n else 0
        op = torch.empty(total, dtype=torch.int32, device=dev)
        ob = torch.empty(total, dtype=torch.int32, device=dev)
        if total:
            N.check(lib.dryad_join_probe(ptr(self.keys), ptr(self.head), self.cap, ptr(self.next),
                                         ptr(probe_keys), n, 1, ptr(offs), ptr(op), ptr(ob), stream()),
                    "join_probe")
            _count()
        return op, ob

    def contains(self, probe_keys: torch.Tensor) -> torch.Tensor:
        require_cuda(probe_keys)
        out = torch.empty(probe_keys.numel(), dtype=torch.uint8, device=probe_keys.device)
        N.check(N.kernels().dryad_set_contains(ptr(self.keys), self.cap, ptr(probe_keys),
                                               probe_keys.numel(), ptr(out), stream()), "set_contains")
        _count()
        return out


def exclusive_scan_i64(x: torch.Tensor):
    """In-place exclusive prefix sum of an int64 column (K16); returns (x, total[1])."""
    require_cuda(x)
    n = x.numel()
    lib = N.kernels()
    total = torch.zeros(1, dtype=torch.int64, device=x.device)
    sc = scratch(lib.dryad_scan_i64_scratch_bytes(n), x.device, "scan64")
    N.check(lib.dryad_exclusive_scan_i64(ptr(x), n, ptr(total), ptr(sc), sc.numel(), stream()), "exclusive_scan_i64")
    _count(3)
    return x, total


# ---- K12 -------------------------------------------------------------------------------------
def compact_flags(flags: torch.Tensor):
    """Ascending indices (int32 bit patterns of u32) of non-zero flags; returns (idx, count)."""
    require_cuda(flags)
    n = flags.numel()
    lib = N.kernels()
    sc = scratch(lib.dryad_compact_scratch_bytes(n), flags.device, "compact")
    idx = torch.empty(n, dtype=torch.int32, device=flags.device)
    cnt = torch.zeros(1, dtype=torch.int64, device=flags.device)
    N.check(lib.dryad_compact_flags(ptr(flags), n, ptr(idx), ptr(cnt), ptr(sc), sc.numel(), stream()),
            "compact_flags")
    _count(3)
    m = int(cnt.item())
    return idx[:m], m


# ---- K7 --------------------------------------------------------------------------------------
def group_by_sort(key_bits: torch.Tensor):
    """Sort-based grouping of a key column (int64 bit patterns).  Returns (order, starts, keys):
    `order` int32[n] is the STABLE sort permutation (rows of one group keep their source order),
    `starts` int32[g] the first position of every group in the sorted order, `keys` int64[g] the
    group keys ascending by bit pattern — a CSR view of the groups."""
    require_cuda(key_bits)
    n = key_bits.numel()
    if n == 0:
        z = torch.empty(0, dtype=torch.int32, device=key_bits.device)
        return z, z.clone(), key_bits[:0]
    skeys, order = radix_sort_pairs(key_bits.clone(), None, iota=True)
    flags = torch.empty(n, dtype=torch.uint8, device=key_bits.device)
    N.check(N.kernels().dryad_mark_boundaries(ptr(skeys), n, ptr(flags), stream()), "mark_boundaries")
    _count()
    starts, g = compact_flags(flags)
    return order, starts, skeys[starts.long()]


# ---- K15 -------------------------------------------------------------------------------------
_W2DT = {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64}


def rows_to_columns(rows: torch.Tensor, fields: Sequence[tuple]):
    """rows: [n, stride] uint8 packed little-endian records; fields: (offset, torch dtype) per
    column.  Returns one tensor per field (the partfile record format -> columnar)."""
    require_cuda(rows)
    n, stride = rows.shape
    nc = len(fields)
    outs = [torch.empty(n, dtype=dt, device=rows.device) for _, dt in fields]
    offs = (C.c_int * nc)(*[int(o) for o, _ in fields])
    widths = (C.c_int * nc)(*[o.element_size() for o in outs])
    ptrs = (C.c_void_p * nc)(*[o.data_ptr() if n else 1 for o in outs])
    N.check(N.kernels().dryad_rows_to_columns(ptr(rows), n, stride, nc, offs, widths, ptrs, stream()), "rows_to_columns")
    _count()
    return outs


def columns_to_rows(cols: Sequence[torch.Tensor], offsets: Sequence[int], stride: int) -> torch.Tensor:
    """Inverse of rows_to_columns: packs columns into [n, stride] uint8 records (gaps are zero)."""
    require_cuda(*cols)
    n = cols[0].numel()
    nc = len(cols)
    rows = torch.empty((n, stride), dtype=torch.uint8, device=cols[0].device)
    offs = (C.c_int * nc)(*[int(o) for o in offsets])
    widths = (C.c_int * nc)(*[c.element_size() for c in cols])
    ptrs = (C.c_void_p * nc)(*[c.data_ptr() if n else 1 for c in cols])
    N.check(N.kernels().dryad_columns_to_rows(ptr(rows), n, stride, nc, offs, widths, ptrs, stream()), "columns_to_rows")
    _count()
    return rows


# ---- K10 -------------------------------------------------------------------------------------
def set_contains(table: "HashAggTable", probe: torch.Tensor) -> torch.Tensor:
    """uint8 flag per probe key: present in the (finished) key table."""
    require_cuda(probe)
    out = torch.empty(probe.numel(), dtype=torch.uint8, device=probe.device)
    N.check(N.kernels().dryad_set_contains(ptr(table.keys), table.cap, ptr(probe), probe.numel(), ptr(out), stream()), "set_contains")
    _count()
    return out


def distinct_keys(key_bits: torch.Tensor) -> torch.Tensor:
    tbl = HashAggTable(agg_capacity(max(key_bits.numel(), 16), 0.5), [], key_bits.device)
    if key_bits.numel():
        tbl.update(key_bits.contiguous(), [])
    return tbl.extract()[0]


def set_operation(op: str, a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Hash-set algebra on key bit patterns: 'union' = Distinct(a ++ b); 'intersect' / 'except' =
    Distinct of the rows of `a` that are (not) present in `b` (LinqToDryad/DryadLinqVertex.cs:1239-1660)."""
    if op == "union":
        return distinct_keys(torch.cat([a, b]))
    if op not in ("intersect", "except"):
        raise ValueError(op)
    if a.numel() == 0:
        return a
    if b.numel() == 0:
        return a[:0] if op == "intersect" else distinct_keys(a)
    tb = HashAggTable(agg_capacity(b.numel(), 0.5), [], b.device)
    tb.update(b.contiguous(), [])
    if int(tb.status.item()) == 1:
        raise N.NativeError("hash set table full")
    flags = set_contains(tb, a.contiguous())
    if op == "except":
        flags = 1 - flags
    idx, _ = compact_flags(flags)
    return distinct_keys(a[idx.long()])


def reduce(col: torch.Tensor, op: str):
    """Deterministic two-stage Sum/Min/Max; returns a 1-element device tensor (int64 or float64)."""
    require_cuda(col)
    lib = N.kernels()
    is_f = col.dtype in (torch.float32, torch.float64)
    out = torch.zeros(1, dtype=torch.float64 if is_f else torch.int64, device=col.device)
    sc = scratch(lib.dryad_reduce_scratch_bytes(), col.device, "reduce")
    if not getattr(reduce, "_zeroed", False):
        sc.zero_()
        reduce._zeroed = True
    N.check(lib.dryad_reduce(ptr(col), DT[col.dtype], col.numel(), RED[op], ptr(out), ptr(sc),
                             sc.numel(), stream()), "reduce")
    _count()
    return out


# ---- K13 -------------------------------------------------------------------------------------
def tokenize(text: torch.Tensor, delim: str = " ", keep_empty: bool = True):
    """text: uint8 CUDA tensor.  Returns (starts int64[m], lens int32[m], hashes int64[m])."""
    require_cuda(text)
    n = text.numel()
    lib = N.kernels()
    sc = scratch(lib.dryad_tokenize_scratch_bytes(n), text.device, "tokenize")
    cnt = torch.zeros(1, dtype=torch.int64, device=text.device)
    N.check(lib.dryad_tokenize_count(ptr(text), n, ord(delim), int(keep_empty), ptr(cnt), ptr(sc),
                                     sc.numel(), stream()), "tokenize_count")
    _count(2)
    m = int(cnt.item())
    starts = torch.empty(m, dtype=torch.int64, device=text.device)
    lens = torch.empty(m, dtype=torch.int32, device=text.device)
    hashes = torch.empty(m, dtype=torch.int64, device=text.device)
    if m:
        N.check(lib.dryad_tokenize_write(ptr(text), n, ord(delim), int(keep_empty), ptr(starts),
                                         ptr(lens), ptr(hashes), ptr(sc), sc.numel(), stream()),
                "tokenize_write")
        _count()
    return starts, lens, hashes


def tokenize_verify(text, starts, lens, rep_start, rep_len) -> int:
    mism = torch.zeros(1, dtype=torch.int64, device=text.device)
    N.check(N.kernels().dryad_tokenize_verify(ptr(text), ptr(starts), ptr(lens), ptr(rep_start),
                                              ptr(rep_len), starts.numel(), ptr(mism), stream()),
            "tokenize_verify")
    _count()
    return int(mism.item())


# ---- K5 / K9 ---------------------------------------------------------------------------------
def merge_pairs(ka: torch.Tensor, va: Optional[torch.Tensor], kb: torch.Tensor, vb: Optional[torch.Tensor]):
    """Stable merge of two ascending (key bits, int32 value) runs (ties: run A first)."""
    require_cuda(ka, kb, va, vb)
    n = ka.numel() + kb.numel()
    ko = torch.empty(n, dtype=torch.int64, device=ka.device)
    vo = torch.empty(n, dtype=torch.int32, device=ka.device) if va is not None else None
    N.check(N.kernels().dryad_merge_pairs_u64(ptr(ka), ptr(va), ka.numel(), ptr(kb), ptr(vb), kb.numel(), ptr(ko), ptr(vo),
                                              stream()), "merge_pairs")
    _count()
    return ko, vo


def kway_merge(runs):
    """Merge k ascending (keys, values) runs by a tree of merge-path 2-way merges; port order
    breaks ties (the reference's MergeSort keeps port order, DryadLinqVertex.cs:319-423)."""
    runs = [r for r in runs]
    if not runs:
        raise ValueError("no runs")
    while len(runs) > 1:
        nxt = []
        for i in range(0, len(runs), 2):
            if i + 1 < len(runs):
                nxt.append(merge_pairs(runs[i][0], runs[i][1], runs[i + 1][0], runs[i + 1][1]))
            else:
                nxt.append(runs[i])
        runs = nxt
    return runs[0]


def merge_join(outer_sorted: torch.Tensor, inner_sorted: torch.Tensor):
    """Join two ASCENDING key columns (unsigned bit order): returns (outer_idx, inner_idx) int32
    pairs in outer order — the MergeJoin of key-sorted inputs."""
    require_cuda(outer_sorted, inner_sorted)
    n, m = outer_sorted.numel(), inner_sorted.numel()
    dev = outer_sorted.device
    lo = torch.empty(n, dtype=torch.int64, device=dev)
    hi = torch.empty(n, dtype=torch.int64, device=dev)
    lib = N.kernels()
    N.check(lib.dryad_sorted_bounds(ptr(inner_sorted), m, ptr(outer_sorted), n, ptr(lo), ptr(hi), stream()), "sorted_bounds")
    _count()
    cnt = hi - lo
    offs = torch.cumsum(cnt, 0) - cnt
    total = int((offs[-1] + cnt[-1]).item()) if n else 0
    oo = torch.empty(total, dtype=torch.int32, device=dev)
    oi = torch.empty(total, dtype=torch.int32, device=dev)
    if total:
        N.check(lib.dryad_emit_ranges(ptr(lo), ptr(hi), ptr(offs), n, ptr(oo), ptr(oi), stream()), "emit_ranges")
        _count()
    return oo, oi
