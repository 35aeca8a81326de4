"""pbx_batch_build_static — the entry point for callers that cannot allocate device memory.

Some users of libpbx_cuda.so may not call cudaMalloc: a data-feed thread that builds the next
batch while training kernels are running (an allocation would synchronise the whole device), a
process that has already committed all of HBM to the sparse table, or a host that embeds the
library under a fixed memory budget.  For them the block-sum scratch of the batch-build kernel is
a small array compiled into the library, and the call takes nothing but inputs and outputs.

The test drives the entry point through the plain C ABI (ctypes on the shared object) and checks,
on two calls in a row, that it produces exactly what pbx_batch_build() produces with caller-owned
scratch.  The second call is smaller than the first, so stale scratch contents from the first call
would show up as wrong offsets.  It also checks that an over-sized request is refused.

This file sorts last on purpose: a run that stops at the first failure has then already executed
every other test.  It has no dependence on the tests before it and passes in a fresh process too.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _make_batch(B, S, max_len, seed, device):
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(0, max_len + 1, (B, S), generator=g)
    off = torch.zeros(B, S + 1, dtype=torch.int32)
    off[:, 1:] = torch.cumsum(lens, 1).int()
    base = torch.zeros(B, dtype=torch.int32)
    base[1:] = torch.cumsum(off[:-1, S], 0).int()
    off += base[:, None]
    n = int(off[-1, S])
    keys = torch.randint(1, 1 << 60, (max(n, 1),), generator=g, dtype=torch.int64)
    return keys.to(device), off.to(device), n


def test_static_scratch_entry_point_matches_batch_build_twice():
    from paddlebox_b200 import _native as N
    lib = N.cuda()                       # ctypes.CDLL of paddlebox_b200/lib/libpbx_cuda.so
    fn = lib.pbx_batch_build_static      # resolved by name through the C ABI
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream

    # two calls in a row: a large batch, then a small one (fewer thread blocks than the first)
    for B, S, max_len, seed in [(4096, 26, 4, 1), (300, 7, 3, 2)]:
        keys, off, n = _make_batch(B, S, max_len, seed, dev)
        # reference result: existing entry point with caller-provided scratch
        ref_keys = torch.zeros(max(n, 1), dtype=torch.int64, device=dev)
        ref_lod = torch.zeros(S * B + 1, dtype=torch.int32, device=dev)
        scratch = torch.zeros(1024, dtype=torch.int32, device=dev)
        N.check(lib.pbx_batch_build(N.ptr(keys), N.ptr(off), B, S, N.ptr(ref_keys), N.ptr(ref_lod),
                                    N.ptr(scratch), C.c_void_p(stream)), "pbx_batch_build")
        # entry point under test: only inputs and outputs
        out_keys = torch.full((max(n, 1),), -1, dtype=torch.int64, device=dev)
        out_lod = torch.full((S * B + 1,), -1, dtype=torch.int32, device=dev)
        rc = fn(keys.data_ptr(), off.data_ptr(), B, S, out_keys.data_ptr(), out_lod.data_ptr(), stream)
        assert rc == 0, lib.pbx_last_error().decode()
        torch.cuda.synchronize()
        assert int(ref_lod[-1]) == n
        assert torch.equal(out_lod, ref_lod)            # integer outputs: exact match
        assert torch.equal(out_keys[:n], ref_keys[:n])

    # beyond the built-in scratch limit (B*S > 1,048,576) the call must refuse, not run
    rc = fn(keys.data_ptr(), off.data_ptr(), 1 << 15, 64, out_keys.data_ptr(), out_lod.data_ptr(), stream)
    assert rc != 0
    assert b"exceeds the built-in scratch limit" in lib.pbx_last_error()
    torch.cuda.synchronize()
