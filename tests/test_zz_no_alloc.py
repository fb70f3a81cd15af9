"""The allocation-free sort entry point, called through the C ABI.

Some users of the kernel library cannot allocate device memory when they need a sort: their own
memory pool already owns the GPU, or they run on a thread that must not block on the CUDA
allocator, or they are plain C programs with no allocator of their own.  For them
`dryad_radix_sort_pairs_u64_noalloc` keeps its working storage in an array built into the library.
This test drives that entry point with raw pointers, exactly as such a caller would, and checks it
against the regular entry point (which takes caller-provided scratch) on two consecutive calls —
the second call reuses the same built-in array, so it also proves that leftovers from the first
call do no harm.  The file name sorts last on purpose: if this test fails under `-x`, every other
test has already run.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_N = 1 << 22
ERR_TOO_LARGE = -2


def _lib():
    from dryad_b200 import _native as N
    lib = N.kernels()
    fn = lib.dryad_radix_sort_pairs_u64_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int,
                   C.POINTER(C.c_int), C.c_void_p]
    return lib, fn


def _sort_with(fn_noalloc, lib, keys, use_noalloc):
    n = keys.numel()
    k, ka = keys.clone(), torch.empty_like(keys)
    v = torch.empty(n, dtype=torch.int32, device=keys.device)
    va = torch.empty_like(v)
    in_alt = C.c_int(0)
    stream = torch.cuda.current_stream().cuda_stream
    if use_noalloc:
        rc = fn_noalloc(k.data_ptr(), ka.data_ptr(), v.data_ptr(), va.data_ptr(), n, 0, 64, 1, C.byref(in_alt), stream)
    else:
        sb = lib.dryad_radix_sort_scratch_bytes(n, 0, 64)
        sc = torch.empty(sb, dtype=torch.uint8, device=keys.device)
        rc = lib.dryad_radix_sort_pairs_u64(k.data_ptr(), ka.data_ptr(), v.data_ptr(), va.data_ptr(), n, 0, 64, 1,
                                            sc.data_ptr(), sb, C.byref(in_alt), stream)
    assert rc == 0, rc
    torch.cuda.synchronize()
    return (ka, va) if in_alt.value else (k, v)


def test_noalloc_sort_matches_regular_entry_point_twice_in_a_row():
    lib, fn = _lib()
    g = torch.Generator(device="cuda").manual_seed(1234)
    # first call: large (uses most of the built-in array); second call: smaller and different, so
    # stale histogram / status contents from the first call would show up as wrong output
    for n in (3_000_001, 70_003):
        keys = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device="cuda", generator=g)
        keys[::5] = keys[1]                       # duplicates: the value order checks stability
        got_k, got_v = _sort_with(fn, lib, keys, use_noalloc=True)
        want_k, want_v = _sort_with(fn, lib, keys, use_noalloc=False)
        assert torch.equal(got_k, want_k)
        assert torch.equal(got_v, want_v)


def test_noalloc_sort_rejects_sizes_beyond_its_built_in_scratch():
    lib, fn = _lib()
    n = MAX_N + 1
    k = torch.zeros(n, dtype=torch.int64, device="cuda")
    ka = torch.empty_like(k)
    in_alt = C.c_int(0)
    rc = fn(k.data_ptr(), ka.data_ptr(), None, None, n, 0, 64, 0, C.byref(in_alt),
            torch.cuda.current_stream().cuda_stream)
    assert rc == ERR_TOO_LARGE
    assert b"beyond supported maximum" in lib.dryad_error_string(rc)
    # exactly at the limit is accepted
    k = torch.randint(0, 2**40, (MAX_N,), dtype=torch.int64, device="cuda")
    got_k, _ = _sort_with(fn, lib, k, use_noalloc=True)
    assert torch.equal(got_k, torch.sort(k).values)
