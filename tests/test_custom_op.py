"""Custom device operators (Operators.device_expr -> NVRTC -> the fused GRS kernel template).

CPU part: the expression compiles for sm_100a without a GPU, diagnostics reach the caller, the
operator still works on the host path through host_fn. GPU part: numerics against a plain PyTorch
evaluation of the same expression in ascending rank (and thread) order - the order the kernels
guarantee, which makes non-commutative expressions well defined and bit-reproducible.
"""
import ctypes as C

import numpy as np
import pytest

from ytk_mp4j_b200 import Mp4jException, Operators
from ytk_mp4j_b200 import _native as N


def _compile(code, expr, world):
    from ytk_mp4j_b200.parallel import gpu_plane  # registers the ctypes prototypes
    n = C.c_longlong()
    rc = N.lib.mp4j_rtc_compile(code, expr.encode(), world, C.byref(n))
    return rc, n.value, (N.lib.mp4j_last_error() or b"").decode()


def _have_nvrtc():
    rc, _, err = _compile(N.F32, "a + b", 2)
    return rc == 0 or "NVRTC" not in err


needs_nvrtc = pytest.mark.skipif(not _have_nvrtc(), reason="libnvrtc not available")


@needs_nvrtc
@pytest.mark.parametrize("code,expr", [(N.F64, "a > b ? a : b"), (N.F32, "a * 0.5f + b"), (N.I64, "a + 3 * b"),
                                       (N.I32, "(a ^ b) + 1"), (N.I16, "a - b"), (N.I8, "a | b")])
@pytest.mark.parametrize("world", [2, 3])
def test_compiles_for_sm100a_without_gpu(code, expr, world):
    rc, nbytes, err = _compile(code, expr, world)
    assert rc == 0, err
    assert nbytes > 10000  # a real SASS image, not PTX


@needs_nvrtc
def test_compile_error_reaches_caller():
    rc, _, err = _compile(N.F32, "a +* b", 4)
    assert rc != 0 and "does not compile" in err and "a +* b" in err
    rc, _, err = _compile(N.F32, "a; }\n#include <x>", 4)
    assert rc != 0 and "may not contain" in err
    rc, _, err = _compile(N.BF16, "a + b", 4)
    assert rc != 0 and "primitive operand types" in err


def test_host_path_uses_host_fn():
    op = Operators.device_expr("a + 2 * b", host_fn=lambda a, b: a + 2 * b)
    assert not op.is_native and op.expr == "a + 2 * b"
    assert op.apply(1.0, 2.0) == 5.0
    with pytest.raises(Mp4jException):
        Operators.device_expr("a + b").apply(1, 2)


# ---------------------------------------------------------------------------------------- GPU
def _fold(parts, fn):
    acc = parts[0].clone()
    for x in parts[1:]:
        acc = fn(acc, x)
    return acc


@pytest.fixture(scope="module")
def planes():
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    cache = {}

    def get(world):
        if world not in cache:
            cache[world] = EmuPlane(world, 0, heap_bytes=128 << 20, staging_bytes=4 << 20)
        return cache[world]
    yield get
    for p in cache.values():
        p.close()


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("n,f", [(7, 0), (4099, 3), ((3 << 20) + 5, 1)])
def test_allreduce_noncommutative_f32(planes, world, n, f):
    import torch
    plane = planes(world)
    op = plane.custom_op("a * 0.5f + b", N.F32)
    assert op >= 64
    g = torch.Generator().manual_seed(n + world)
    host = [torch.randn(n, generator=g) for _ in range(world)]
    dev = [h.cuda() for h in host]
    plane.allreduce(dev, op, f, n)
    torch.cuda.synchronize()
    plane.check()
    want = _fold([h[f:].cuda() for h in host], lambda a, b: a * 0.5 + b)
    for r in range(world):
        assert torch.equal(dev[r][f:], want), (world, n, r, plane.last())
        assert torch.equal(dev[r][:f].cpu(), host[r][:f])


@pytest.mark.gpu
@pytest.mark.parametrize("code,expr,fn", [
    (N.F64, "a > b ? a : b", lambda a, b: __import__("torch").maximum(a, b)),
    (N.I32, "(a ^ b) + 1", lambda a, b: (a ^ b) + 1),
    (N.I64, "a + 3 * b", lambda a, b: a + 3 * b),
    (N.I16, "a - b", lambda a, b: a - b),
    (N.I8, "a | b", lambda a, b: a | b),
])
def test_allreduce_other_types_symmetric(planes, code, expr, fn):
    import torch
    tdt = {N.F64: torch.float64, N.I32: torch.int32, N.I64: torch.int64, N.I16: torch.int16, N.I8: torch.int8}[code]
    world, n = 4, 100003
    plane = planes(world)
    op = plane.custom_op(expr, code)
    g = torch.Generator().manual_seed(code)
    host = [torch.randn(n, generator=g, dtype=tdt) if tdt.is_floating_point else
            torch.randint(-50, 50, (n,), generator=g, dtype=tdt) for _ in range(world)]
    dev = plane.empty(n, code)  # symmetric: zero-copy path
    for d, h in zip(dev, host):
        d.copy_(h)
    plane.allreduce(dev, op, 0, n)
    torch.cuda.synchronize()
    want = _fold([h.cuda() for h in host], fn)
    for r in range(world):
        assert torch.equal(dev[r], want), (expr, r)
    assert "staged=0" in plane.last()


@pytest.mark.gpu
def test_reduce_scatter_and_reduce(planes):
    import torch
    world, n = 4, 50001
    plane = planes(world)
    op = plane.custom_op("a - 2 * b", N.I32)
    g = torch.Generator().manual_seed(5)
    host = [torch.randint(-1000, 1000, (n,), generator=g, dtype=torch.int32) for _ in range(world)]
    want = _fold(host, lambda a, b: a - 2 * b)
    counts = [n // world] * (world - 1) + [n - (n // world) * (world - 1) - 1]
    dev = [h.cuda() for h in host]
    plane.reduce_scatter(dev, op, 1, counts)
    torch.cuda.synchronize()
    lo = 1
    for r in range(world):
        assert torch.equal(dev[r][lo:lo + counts[r]].cpu(), want[lo:lo + counts[r]]), r
        lo += counts[r]
    dev = [h.cuda() for h in host]
    plane.reduce(dev, op, 0, n, 2)
    torch.cuda.synchronize()
    assert torch.equal(dev[2].cpu(), want)


@pytest.mark.gpu
@pytest.mark.parametrize("fanout", [True, False])
def test_thread_allreduce_custom(planes, fanout):
    import torch
    world, T, n = 2, 3, 70001
    plane = planes(world)
    op = plane.custom_op("a * 0.5f + b", N.F32)
    g = torch.Generator().manual_seed(11)
    host = [[torch.randn(n, generator=g) for _ in range(T)] for _ in range(world)]
    dev = [[h.cuda() for h in row] for row in host]
    dst = 0 if fanout else 1
    plane.thread_allreduce(dev, op, 0, n, fanout=fanout, dst=dst, local_only=not fanout)
    torch.cuda.synchronize()
    fn = lambda a, b: a * 0.5 + b
    if fanout:
        per_rank = [_fold([row[0]] + row[1:], fn) for row in host]
        want = _fold(per_rank, fn).cuda()
        for r in range(world):
            for t in range(T):
                assert torch.equal(dev[r][t], want), (r, t)
    else:
        for r in range(world):
            order = [host[r][dst]] + [host[r][t] for t in range(T) if t != dst]
            assert torch.equal(dev[r][dst].cpu(), _fold(order, fn)), r


@pytest.mark.gpu
def test_custom_matches_predefined_sum_bitwise(planes):
    """`a + b` through NVRTC is the same arithmetic, in the same order, as the built-in SUM."""
    import torch
    world, n = 8, (1 << 20) + 7
    plane = planes(world)
    op = plane.custom_op("a + b", N.F32)
    g = torch.Generator().manual_seed(3)
    host = [torch.randn(n, generator=g) for _ in range(world)]
    a = [h.cuda() for h in host]
    b = [h.cuda() for h in host]
    plane.allreduce(a, op, 0, n)
    plane.allreduce(b, N.SUM, 0, n, algo=N.ALGO_TWOSHOT)
    torch.cuda.synchronize()
    assert torch.equal(a[0], b[0]) and torch.equal(a[7], b[7])
