"""GPU numerics: every dense collective kernel against a plain PyTorch reference of the same op.

Multi-rank logic runs on ONE GPU through EmuPlane (all ranks' rows in one cooperative launch), as
the B200 profiling guide prescribes for flag-waiting kernels. Float SUM is compared bit-exactly:
the kernels reduce in ascending rank order, so a sequential fp32 sum is the same arithmetic.
Value patterns follow the reference's check harness plus what it never covers:
non-zero roots, from > 0, odd sizes, empty segments, every operator.
"""
import pytest
import torch

from ytk_mp4j_b200 import _native as N

pytestmark = pytest.mark.gpu

DTYPES = {N.F64: torch.float64, N.F32: torch.float32, N.I64: torch.int64, N.I32: torch.int32,
          N.I16: torch.int16, N.I8: torch.int8}


@pytest.fixture(scope="module")
def planes():
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    cache = {}

    def get(world):
        if world not in cache:
            cache[world] = EmuPlane(world, 0, heap_bytes=192 << 20, staging_bytes=8 << 20)
        return cache[world]
    yield get
    for p in cache.values():
        p.close()


def make_inputs(plane, world, n, dtype, symmetric, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    if dtype.is_floating_point:
        host = [torch.randn(n, generator=g, dtype=dtype) for _ in range(world)]
    else:
        info = torch.iinfo(dtype)
        lo, hi = max(info.min, -100), min(info.max, 100)
        host = [torch.randint(lo, hi, (n,), generator=g, dtype=dtype) for _ in range(world)]
    if symmetric:
        code = {v: k for k, v in DTYPES.items()}[dtype]
        dev = plane.empty(n, code)
        for d, h in zip(dev, host):
            d.copy_(h)
    else:
        dev = [h.cuda() for h in host]
    return host, dev


def seq_reduce(parts, op):
    acc = parts[0].clone()
    for x in parts[1:]:
        if op == N.SUM:
            acc = acc + x
        elif op == N.MAX:
            acc = torch.maximum(acc, x)
        elif op == N.MIN:
            acc = torch.minimum(acc, x)
        elif op == N.PROD:
            acc = acc * x
        elif op == N.BAND:
            acc = acc & x
        elif op == N.BOR:
            acc = acc | x
        elif op == N.BXOR:
            acc = acc ^ x
    return acc


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("n,f,algo,symmetric", [
    (1, 0, N.ALGO_AUTO, False),
    (1000, 3, N.ALGO_AUTO, False),         # one-shot, unaligned from
    (1000, 3, N.ALGO_TWOSHOT, False),      # two-shot staged, unaligned from
    (70001, 0, N.ALGO_TWOSHOT, True),      # two-shot zero-copy, odd size
    (70001, 5, N.ALGO_ONESHOT, True),
    ((1 << 20) + 5, 1, N.ALGO_AUTO, True),
    ((3 << 20) + 1, 0, N.ALGO_AUTO, False),  # staged and windowed (> staging window)
])
def test_allreduce_f32_sum_exact(planes, world, n, f, algo, symmetric):
    plane = planes(world)
    host, dev = make_inputs(plane, world, n, torch.float32, symmetric, 1234 + n)
    plane.allreduce(dev, N.SUM, f, n, algo)
    torch.cuda.synchronize()
    plane.check()
    want = seq_reduce([h[f:] for h in host], N.SUM)
    for r in range(world):
        assert torch.equal(dev[r][f:].cpu(), want), (r, plane.last())
        assert torch.equal(dev[r][:f].cpu(), host[r][:f]), "elements before `from` must be untouched"


@pytest.mark.parametrize("code", list(DTYPES))
def test_allreduce_all_dtypes_all_ops(planes, code):
    world, n = 4, 4099
    plane = planes(world)
    dtype = DTYPES[code]
    ops = [N.SUM, N.MAX, N.MIN, N.PROD]
    if not dtype.is_floating_point:
        ops += [N.BAND, N.BOR, N.BXOR]
    for op in ops:
        for algo in (N.ALGO_ONESHOT, N.ALGO_TWOSHOT):
            host, dev = make_inputs(plane, world, n, dtype, algo == N.ALGO_TWOSHOT, 7 + op)
            if op == N.PROD:  # keep products finite / wrap-around still defined
                host = [(h.sign() if dtype.is_floating_point else (h % 3 - 1).to(dtype)) for h in host]
                for d, h in zip(dev, host):
                    d.copy_(h)
            plane.allreduce(dev, op, 0, n, algo)
            torch.cuda.synchronize()
            want = seq_reduce(host, op)
            for r in range(world):
                assert torch.equal(dev[r].cpu(), want), (code, op, algo, r)
    plane.check()


def test_float_max_min_nan_and_signed_zero(planes):
    world = 2
    plane = planes(world)
    a = torch.tensor([float("nan"), 1.0, -0.0, 0.0, 5.0], dtype=torch.float32)
    b = torch.tensor([1.0, float("nan"), 0.0, -0.0, 7.0], dtype=torch.float32)
    dev = [a.cuda(), b.cuda()]
    plane.allreduce(dev, N.MAX, 0, 5, N.ALGO_ONESHOT)
    out = dev[0].cpu()
    assert torch.isnan(out[0]) and torch.isnan(out[1])  # Math.max propagates NaN
    assert not torch.signbit(out[2]) and not torch.signbit(out[3]) and out[4] == 7
    dev = [a.cuda(), b.cuda()]
    plane.allreduce(dev, N.MIN, 0, 5, N.ALGO_ONESHOT)
    out = dev[1].cpu()
    assert torch.isnan(out[0]) and torch.isnan(out[1])
    assert torch.signbit(out[2]) and torch.signbit(out[3]) and out[4] == 5


def test_loc_operators_on_gpu(planes):
    from ytk_mp4j_b200 import Operators
    world = 4
    plane = planes(world)
    D, L = Operators.Double, Operators.Long
    dev = [torch.tensor([D.compositeDouble(float(r), r), D.compositeDouble(1.0, r), D.compositeDouble(-float(r), r)],
                        dtype=torch.float64).cuda() for r in range(world)]
    plane.allreduce(dev, N.MAXLOC, 0, 3, N.ALGO_ONESHOT)
    out = dev[2].cpu().tolist()
    assert D.getIntLoc(out[0]) == world - 1 and D.getFloatVal(out[0]) == world - 1
    assert D.getIntLoc(out[1]) == 0  # tie -> lowest loc
    assert D.getIntLoc(out[2]) == 0
    dev = [torch.tensor([L.compositeLong(r, r), L.compositeLong(9, r)], dtype=torch.int64).cuda()
           for r in range(world)]
    plane.allreduce(dev, N.MINLOC, 0, 2, N.ALGO_TWOSHOT)
    out = dev[1].cpu().tolist()
    assert L.getIntLoc(out[0]) == 0 and L.getIntLoc(out[1]) == 0 and L.getIntVal(out[1]) == 9


@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("symmetric", [False, True])
def test_reduce_scatter_irregular_counts(planes, world, symmetric):
    plane = planes(world)
    counts = [(37 * (q + 1)) % 11 * 100 + (0 if q == 1 else q) for q in range(world)]  # includes zeros/odd
    f = 5
    n = f + sum(counts) + 9
    host, dev = make_inputs(plane, world, n, torch.float64, symmetric, 99)
    plane.reduce_scatter(dev, N.SUM, f, counts)
    torch.cuda.synchronize()
    want = seq_reduce(host, N.SUM)
    off = f
    for r in range(world):
        got = dev[r].cpu()
        assert torch.equal(got[off:off + counts[r]], want[off:off + counts[r]]), r
        assert torch.equal(got[:f], host[r][:f]) and torch.equal(got[f + sum(counts):], host[r][f + sum(counts):])
        off += counts[r]
    plane.check()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("code", [N.F32, N.I8, N.I64])
def test_allgather_gather_scatter_broadcast_reduce(planes, world, code):
    plane = planes(world)
    dtype = DTYPES[code]
    n = 10007
    avg = n // world
    froms = [r * avg for r in range(world)]
    tos = [(r + 1) * avg if r < world - 1 else n for r in range(world)]
    owner = torch.empty(n, dtype=dtype)
    for r in range(world):
        owner[froms[r]:tos[r]] = r
    root = world - 1
    # allgather: own slice = rank (ProcessAllgatherCheck.java:55-83)
    dev = []
    for r in range(world):
        t = torch.full((n,), -1, dtype=dtype)
        t[froms[r]:tos[r]] = r
        dev.append(t.cuda())
    plane.allgather(dev, froms, tos)
    for r in range(world):
        assert torch.equal(dev[r].cpu(), owner), ("allgather", r)
    # gather to a non-zero root (ProcessGatherCheck.java:54-84 only tests root 0)
    dev = []
    for r in range(world):
        t = torch.full((n,), -1, dtype=dtype)
        t[froms[r]:tos[r]] = r
        dev.append(t.cuda())
    plane.gather(dev, froms, tos, root)
    assert torch.equal(dev[root].cpu(), owner), "gather"
    # scatter (ProcessScatterCheck.java:55-90)
    dev = [(owner.clone() if r == root else torch.full((n,), -1, dtype=dtype)).cuda() for r in range(world)]
    plane.scatter(dev, froms, tos, root)
    for r in range(world):
        got = dev[r].cpu()
        assert torch.all(got[froms[r]:tos[r]] == r), ("scatter", r)
    # broadcast with from > 0 (ProcessBroadcastCheck.java:59-77)
    dev = [torch.full((n,), 1 if r == root else -1, dtype=dtype).cuda() for r in range(world)]
    plane.broadcast(dev, 3, n, root)
    for r in range(world):
        got = dev[r].cpu()
        assert torch.all(got[3:] == 1), ("broadcast", r)
        if r != root:
            assert torch.all(got[:3] == -1)
    # reduce to root (ProcessReduceCheck.java:58-75)
    dev = [torch.ones(n, dtype=dtype).cuda() for _ in range(world)]
    plane.reduce(dev, N.SUM, 0, n, root)
    assert torch.all(dev[root].cpu() == world), "reduce"
    torch.cuda.synchronize()
    plane.check()


def test_zero_copy_symmetric_tensors_report_unstaged(planes):
    plane = planes(4)
    dev = plane.empty(1 << 18, N.F32)
    for r, d in enumerate(dev):
        d.fill_(r + 1)
    plane.allreduce(dev, N.SUM, 0, 1 << 18, N.ALGO_TWOSHOT)
    torch.cuda.synchronize()
    assert "staged=0" in plane.last() and "twoshot" in plane.last()
    assert torch.all(dev[3] == 10)


def test_invalid_operator_raises(planes):
    from ytk_mp4j_b200 import Mp4jException
    plane = planes(2)
    dev = [torch.zeros(8, dtype=torch.float32).cuda() for _ in range(2)]
    with pytest.raises(Mp4jException):
        plane.allreduce(dev, N.BAND, 0, 8)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_staged_windows_pack_one_subrange_per_rank(world):
    """Ordinary (non-symmetric) tensors larger than the staging window: every round carries sub-range j of EVERY
    rank's segment (gpu_plane.cu run_grs). Ragged segments, from > 0, sizes that do not divide."""
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    plane = EmuPlane(world, 0, heap_bytes=160 << 20, staging_bytes=1 << 20)
    try:
        n = 1500007  # 6 MB of fp32 through a 1 MiB window
        g = torch.Generator().manual_seed(world)
        host = [torch.randn(n, generator=g) for _ in range(world)]
        total = host[0].clone()
        for h in host[1:]:
            total = total + h
        cuts = sorted(torch.randint(5, n - 5, (world - 1,), generator=g).tolist())
        froms, tos = [5] + cuts, cuts + [n - 3]
        counts = [t - f for f, t in zip(froms, tos)]
        dev = [h.cuda() for h in host]
        plane.allreduce(dev, N.SUM, 7, n - 1, algo=N.ALGO_TWOSHOT)
        torch.cuda.synchronize()
        for r in range(world):
            got = dev[r].cpu()
            assert torch.equal(got[7:n - 1], total[7:n - 1]) and torch.equal(got[:7], host[r][:7]), r
        dev = [h.cuda() for h in host]
        plane.reduce_scatter(dev, N.SUM, 5, counts, algo=N.ALGO_TWOSHOT)
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(dev[r].cpu()[froms[r]:tos[r]], total[froms[r]:tos[r]]), r
        dev = [h.cuda() for h in host]
        plane.allgather(dev, froms, tos, algo=N.ALGO_TWOSHOT)
        torch.cuda.synchronize()
        for r in range(world):
            got = dev[r].cpu()
            for q in range(world):
                assert torch.equal(got[froms[q]:tos[q]], host[q][froms[q]:tos[q]]), (r, q)
            assert torch.equal(got[:5], host[r][:5]) and torch.equal(got[n - 3:], host[r][n - 3:])
        plane.check()
        assert "staged=1" in plane.last()
    finally:
        plane.close()
