"""Host-side twins of the device hash functions (csrc/kernels/common.cuh), so the CPU oracle, the
planner and the GPU kernels agree on every bucket decision.

Bucket rule = the reference's HashPartition rule, `(hash & 0x7FFFFFFF) % n`
(LinqToDryad/DryadLinqVertex.cs:4861-4862).
"""
from __future__ import annotations

import struct

import numpy as np

M64 = (1 << 64) - 1


def mix64(x: int) -> int:
    x &= M64
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & M64
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & M64
    x ^= x >> 31
    return x


def mix64_np(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(30)
        x *= np.uint64(0xBF58476D1CE4E5B9)
        x ^= x >> np.uint64(27)
        x *= np.uint64(0x94D049BB133111EB)
        x ^= x >> np.uint64(31)
    return x


def hash_bytes_py(b: bytes) -> int:
    h = 0xCBF29CE484222325
    for c in b:
        h ^= c
        h = (h * 0x100000001B3) & M64
    return mix64(h)


def key_bits(v) -> int:
    """Canonical 64-bit pattern of a scalar key (what the device sees in the key column)."""
    if isinstance(v, bool):
        return int(v)
    if isinstance(v, (int, np.integer)):
        return int(v) & M64
    if isinstance(v, (float, np.floating)):
        return struct.unpack("<Q", struct.pack("<d", float(v)))[0]
    if isinstance(v, str):
        return hash_bytes_py(v.encode("utf-8"))
    if isinstance(v, (bytes, bytearray)):
        return hash_bytes_py(bytes(v))
    if isinstance(v, tuple):
        h = 0x9E3779B97F4A7C15
        for e in v:
            h = mix64(h ^ key_bits(e))
        return h
    if v is None:
        return 0
    return hash_bytes_py(repr(v).encode("utf-8"))


def hash_bucket_bits(bits: int, nparts: int) -> int:
    h = mix64(bits)
    h32 = (h ^ (h >> 32)) & 0xFFFFFFFF
    return (h32 & 0x7FFFFFFF) % nparts


def hash_bucket(v, nparts: int) -> int:
    return hash_bucket_bits(key_bits(v), nparts)


def hash_bucket_np(bits: np.ndarray, nparts: int) -> np.ndarray:
    h = mix64_np(bits)
    h32 = (h ^ (h >> np.uint64(32))) & np.uint64(0xFFFFFFFF)
    return ((h32 & np.uint64(0x7FFFFFFF)) % np.uint64(nparts)).astype(np.int64)
