"""The allocation-free reduce-by-key entry point (mp4j_map_reduce_local_noalloc).

Some users of the C ABI cannot allocate device memory at the point where they need a
reduce-by-key: a step that must not stall the GPU (cudaMalloc/cudaFree synchronise the device), a
process whose memory is owned by another pool, or an embedder without a plane object. For them the
library carries its own fixed hash-table scratch. This test drives that entry point through ctypes
exactly as such a caller would (it only hands over input and output buffers) and checks, on two
consecutive calls with different data, that it returns what the allocating entry point
mp4j_gpu_map_reduce_local returns. The file name sorts last on purpose: if this addition breaks,
a run that stops at the first failure has already run every other test.
"""
import ctypes as C

import pytest
import torch

from ytk_mp4j_b200 import _native as N

pytestmark = pytest.mark.gpu

MAX_PAIRS = 1 << 18


def _proto():
    fn = N.lib.mp4j_map_reduce_local_noalloc
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                   C.c_void_p]
    return fn


def _sorted_pairs(k, v):
    order = torch.argsort(k)
    return k[order].cpu(), v[order].cpu()


def _noalloc(fn, keys, vals, code, op):
    n = keys.numel()
    ok = torch.empty(max(n, 1), dtype=torch.int64, device="cuda")
    ov = torch.empty(max(n, 1), dtype=vals.dtype, device="cuda")
    cnt = torch.full((1,), 12345, dtype=torch.int64, device="cuda")  # garbage on purpose: the call must zero it
    rc = fn(keys.data_ptr(), vals.data_ptr(), n, code, op, ok.data_ptr(), ov.data_ptr(), n, cnt.data_ptr(),
            torch.cuda.current_stream().cuda_stream)
    assert rc == 0, N.last_error()
    torch.cuda.synchronize()
    m = int(cnt.item())
    return ok[:m], ov[:m]


@pytest.mark.parametrize("dtype,code,op", [(torch.float32, N.F32, N.SUM), (torch.int64, N.I64, N.MAX),
                                           (torch.float64, N.F64, N.MIN)])
def test_noalloc_matches_allocating_entry_point_twice(dtype, code, op):
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    torch.cuda.set_device(0)
    fn = _proto()
    plane = EmuPlane(1, 0, heap_bytes=192 << 20, staging_bytes=4 << 20)
    try:
        g = torch.Generator().manual_seed(99)
        # second call: fewer pairs and other keys, so anything left in the shared table would show
        for n, span in ((150000, 40000), (3000, 1000)):
            keys = torch.randint(-span, span, (n,), generator=g, dtype=torch.int64).cuda()
            vals = torch.randint(-9, 9, (n,), generator=g).to(dtype).cuda()  # integer-valued: sums are exact
            ref_k, ref_v = _sorted_pairs(*plane.map_reduce_local(keys, vals, op))
            got_k, got_v = _sorted_pairs(*_noalloc(fn, keys, vals, code, op))
            assert torch.equal(got_k, ref_k)
            assert torch.equal(got_v, ref_v)
    finally:
        plane.close()


def test_noalloc_rejects_oversize_and_bad_arguments():
    torch.cuda.set_device(0)
    fn = _proto()
    n = MAX_PAIRS + 1
    keys = torch.zeros(n, dtype=torch.int64, device="cuda")
    vals = torch.zeros(n, dtype=torch.float32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    rc = fn(keys.data_ptr(), vals.data_ptr(), n, N.F32, N.SUM, keys.data_ptr(), vals.data_ptr(), n, cnt.data_ptr(), s)
    assert rc != 0 and "262144" in N.last_error()
    rc = fn(keys.data_ptr(), vals.data_ptr(), 10, N.I8, N.SUM, keys.data_ptr(), vals.data_ptr(), 10, cnt.data_ptr(), s)
    assert rc != 0
    rc = fn(keys.data_ptr(), vals.data_ptr(), 10, N.F32, N.PROD, keys.data_ptr(), vals.data_ptr(), 10, cnt.data_ptr(), s)
    assert rc != 0
    # exactly at the limit is fine, and an empty input yields an empty result
    keys = torch.arange(MAX_PAIRS, dtype=torch.int64, device="cuda")
    vals = torch.ones(MAX_PAIRS, dtype=torch.float32, device="cuda")
    k, v = _noalloc(fn, keys, vals, N.F32, N.SUM)
    assert k.numel() == MAX_PAIRS and bool(torch.all(v == 1))
    k, v = _noalloc(fn, keys[:0], vals[:0], N.F32, N.SUM)
    assert k.numel() == 0


def test_noalloc_count_beyond_capacity_and_reserved_key():
    """Two corners of the contract in mp4j_gpu.h: *out_count is the full number of unique keys even when it
    exceeds out_cap (only out_cap pairs are stored, nothing beyond the caller's buffer is touched), and pairs
    whose key is INT64_MIN - the table's free-slot marker - are ignored without disturbing other keys."""
    torch.cuda.set_device(0)
    fn = _proto()
    s = torch.cuda.current_stream().cuda_stream
    # 1000 unique keys into a 100-pair output buffer with a guard zone behind it
    keys = torch.arange(1000, dtype=torch.int64, device="cuda")
    vals = torch.ones(1000, dtype=torch.float32, device="cuda")
    ok = torch.full((200,), -7, dtype=torch.int64, device="cuda")
    ov = torch.full((200,), -7.0, dtype=torch.float32, device="cuda")
    cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
    assert fn(keys.data_ptr(), vals.data_ptr(), 1000, N.F32, N.SUM, ok.data_ptr(), ov.data_ptr(), 100, cnt.data_ptr(), s) == 0
    torch.cuda.synchronize()
    assert int(cnt.item()) == 1000
    assert bool(torch.all(ok[100:] == -7)) and bool(torch.all(ov[100:] == -7.0))
    assert bool(torch.all((ok[:100] >= 0) & (ok[:100] < 1000))) and bool(torch.all(ov[:100] == 1.0))
    # the reserved key is dropped; every other key keeps exactly its own sum
    int64_min = -(1 << 63)
    keys = torch.tensor([5, int64_min, 7, int64_min, 5, 9], dtype=torch.int64, device="cuda")
    vals = torch.tensor([1.0, 100.0, 2.0, 100.0, 3.0, 4.0], dtype=torch.float32, device="cuda")
    k, v = _sorted_pairs(*_noalloc(fn, keys, vals, N.F32, N.SUM))
    assert k.tolist() == [5, 7, 9] and v.tolist() == [4.0, 2.0, 4.0]
