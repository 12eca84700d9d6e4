"""ThreadCommSlave gather / allgather / broadcast / scatter / reduce-scatter / reduce on CUDA tensors as ONE
launch (csrc/gpu/thread_seg.cu), on emulated ranks, against sequential PyTorch in the reference's order
(thread-id order inside a rank, then rank order: TCS:448, 621, 802, 1010, 1174, 1396). fp32 sums are compared
bit for bit; everything outside a destination's slices must stay untouched."""
import random

import pytest
import torch

from ytk_mp4j_b200 import _native as N

pytestmark = pytest.mark.gpu


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


def _make(world, T, n, dtype=torch.float32, seed=0):
    g = torch.Generator().manual_seed(seed + world * 100 + T)
    if dtype.is_floating_point:
        host = [[torch.randn(n, generator=g, dtype=dtype) for _ in range(T)] for _ in range(world)]
    else:
        host = [[torch.randint(-100, 100, (n,), generator=g, dtype=dtype) for _ in range(T)] for _ in range(world)]
    return host, [[h.cuda() for h in row] for row in host]


def _total(host, op=N.SUM):
    f = (lambda a, b: a + b) if op == N.SUM else torch.maximum
    per_rank = []
    for row in host:
        acc = row[0].clone()
        for x in row[1:]:
            acc = f(acc, x)
        per_rank.append(acc)
    acc = per_rank[0].clone()
    for x in per_rank[1:]:
        acc = f(acc, x)
    return acc


def _ragged(world, T, n, f, seed):
    """[world][T] slice lengths summing to <= n - f, with some empty slices and odd (unaligned) sizes."""
    rnd = random.Random(seed)
    w = [[rnd.choice([0, 1, 3, 5, 8, 13]) for _ in range(T)] for _ in range(world)]
    w[rnd.randrange(world)][rnd.randrange(T)] += 1
    tot = sum(sum(r) for r in w)
    room = n - f - 7
    counts = [[room * x // tot for x in r] for r in w]
    return counts


WORLDS_T = [(1, 2), (1, 8), (2, 3), (2, 8), (4, 2), (4, 3), (8, 2), (8, 8)]


@pytest.mark.parametrize("world,T", WORLDS_T)
@pytest.mark.parametrize("n,f", [(70001, 0), (300007, 1234), (2500000, 5)])
def test_thread_reduce_scatter(planes, world, T, n, f):
    plane = planes(world)
    host, dev = _make(world, T, n, seed=n)
    counts = _ragged(world, T, n, f, n + T)
    segs, off = [], f
    for q in range(world):
        for t in range(T):
            segs.append((off, off + counts[q][t], -1, -1, q, t))
            off += counts[q][t]
    plane.thread_segments("thread_reduce_scatter", dev, N.SUM, segs)
    torch.cuda.synchronize()
    plane.check()
    assert "fused-segments" in plane.last()
    total = _total(host)
    for (lo, hi, _, _, q, t) in segs:
        got = dev[q][t].cpu()
        assert torch.equal(got[lo:hi], total[lo:hi]), (q, t, lo, hi)
        assert torch.equal(got[:lo], host[q][t][:lo]) and torch.equal(got[hi:], host[q][t][hi:]), (q, t, "outside")


@pytest.mark.parametrize("world,T,root", [(1, 3, (0, 2)), (2, 8, (1, 5)), (4, 3, (2, 0)), (8, 2, (7, 1))])
@pytest.mark.parametrize("dtype,op", [(torch.float32, N.SUM), (torch.float64, N.MAX), (torch.int32, N.SUM)])
def test_thread_reduce_to_root_thread(planes, world, T, root, dtype, op):
    plane = planes(world)
    n, f, to = 200003, 17, 199990
    host, dev = _make(world, T, n, dtype, seed=3)
    plane.thread_segments("thread_reduce", dev, op, [(f, to, -1, -1, root[0], root[1])])
    torch.cuda.synchronize()
    plane.check()
    total = _total(host, op)
    for q in range(world):
        for t in range(T):
            got = dev[q][t].cpu()
            if (q, t) == root:
                assert torch.equal(got[f:to], total[f:to])
                assert torch.equal(got[:f], host[q][t][:f]) and torch.equal(got[to:], host[q][t][to:])
            else:
                assert torch.equal(got, host[q][t]), (q, t)


@pytest.mark.parametrize("world,T", WORLDS_T)
def test_thread_allgather_and_gather(planes, world, T):
    plane = planes(world)
    n, f = 250001, 3
    counts = _ragged(world, T, n, f, 11)
    ranges, off = {}, f
    for q in range(world):
        for t in range(T):
            ranges[(q, t)] = (off, off + counts[q][t])
            off += counts[q][t]
    # allgather: everybody ends with every (q, t) slice
    host, dev = _make(world, T, n, seed=5)
    plane.thread_segments("thread_allgather", dev, N.COPY, [(lo, hi, q, t, -1, -1) for (q, t), (lo, hi) in ranges.items()])
    torch.cuda.synchronize()
    plane.check()
    want = host[0][0].clone()
    for (q, t), (lo, hi) in ranges.items():
        want[lo:hi] = host[q][t][lo:hi]
    for q in range(world):
        for t in range(T):
            got = dev[q][t].cpu()
            assert torch.equal(got[f:off], want[f:off]), (q, t)
            assert torch.equal(got[:f], host[q][t][:f]) and torch.equal(got[off:], host[q][t][off:])
    # gather to a non-zero root thread: only the root's array changes
    root = (world - 1, T - 1)
    host, dev = _make(world, T, n, seed=6)
    plane.thread_segments("thread_gather", dev, N.COPY,
                          [(lo, hi, q, t, root[0], root[1]) for (q, t), (lo, hi) in ranges.items()])
    torch.cuda.synchronize()
    plane.check()
    want = host[root[0]][root[1]].clone()
    for (q, t), (lo, hi) in ranges.items():
        want[lo:hi] = host[q][t][lo:hi]
    for q in range(world):
        for t in range(T):
            assert torch.equal(dev[q][t].cpu(), want if (q, t) == root else host[q][t]), (q, t)


@pytest.mark.parametrize("world,T,root", [(1, 2, (0, 1)), (2, 3, (1, 2)), (4, 8, (3, 4)), (8, 2, (5, 0))])
@pytest.mark.parametrize("dtype", [torch.float32, torch.int8, torch.bfloat16])
def test_thread_scatter_and_broadcast(planes, world, T, root, dtype):
    plane = planes(world)
    n, f = 180003, 9
    counts = _ragged(world, T, n, f, 23)
    ranges, off = {}, f
    for q in range(world):
        for t in range(T):
            ranges[(q, t)] = (off, off + counts[q][t])
            off += counts[q][t]
    mk = (lambda seed: _make(world, T, n, dtype, seed)) if dtype != torch.bfloat16 else \
        (lambda seed: (lambda hd: ([[h.bfloat16() for h in r] for r in hd[0]], [[d.bfloat16() for d in r] for r in hd[1]]))(
            _make(world, T, n, torch.float32, seed)))
    host, dev = mk(7)
    plane.thread_segments("thread_scatter", dev, N.COPY,
                          [(lo, hi, root[0], root[1], q, t) for (q, t), (lo, hi) in ranges.items()])
    torch.cuda.synchronize()
    plane.check()
    src = host[root[0]][root[1]]
    for (q, t), (lo, hi) in ranges.items():
        want = host[q][t].clone()
        want[lo:hi] = src[lo:hi]
        assert torch.equal(dev[q][t].cpu(), want), (q, t)
    host, dev = mk(8)
    to = n - 11
    plane.thread_segments("thread_broadcast", dev, N.COPY, [(f, to, root[0], root[1], -1, -1)])
    torch.cuda.synchronize()
    plane.check()
    src = host[root[0]][root[1]]
    for q in range(world):
        for t in range(T):
            want = host[q][t].clone()
            want[f:to] = src[f:to]
            assert torch.equal(dev[q][t].cpu(), want), (q, t)


def test_thread_segments_misaligned_views(planes):
    """Thread buffers that are views at different 4-byte offsets: no common 16-byte alignment, all scalar."""
    plane = planes(2)
    T, n = 3, 50000
    g = torch.Generator().manual_seed(1)
    base = [[torch.randn(n + 8, generator=g).cuda() for _ in range(T)] for _ in range(2)]
    dev = [[base[q][t][(q + t) % 4:(q + t) % 4 + n] for t in range(T)] for q in range(2)]
    host = [[d.cpu().clone() for d in row] for row in dev]
    plane.thread_segments("thread_reduce", dev, N.SUM, [(0, n, -1, -1, 1, 2)])
    torch.cuda.synchronize()
    plane.check()
    assert torch.equal(dev[1][2].cpu(), _total(host))


@pytest.mark.parametrize("world,T,root", [(4, 3, (2, 1)), (8, 2, (5, 0)), (8, 4, (0, 3))])
def test_rooted_collectives_relay_through_all_ranks(planes, world, T, root):
    """broadcastArray / reduceArray beyond 2 ranks and beyond seg_relay_min_bytes run as two launches (scatter +
    allgather, reduce-scatter + gather) so that no single NVLink port carries (p - 1) x the array. Same results,
    bit for bit; like the reference's reduce (reduce-scatter + gather in place, PCS:1390-1416) the relay uses
    [from, to) of thread 0 of the non-root ranks as scratch."""
    plane = planes(world)
    plane.set_option("seg_relay_min_bytes", 0)
    try:
        n, f, to = 600011, 13, 599990
        host, dev = _make(world, T, n, torch.float32, seed=11)
        plane.thread_segments("thread_broadcast", dev, N.COPY, [(f, to, root[0], root[1], -1, -1)])
        torch.cuda.synchronize()
        plane.check()
        assert "relay" in plane.last(), plane.last()
        src = host[root[0]][root[1]]
        for q in range(world):
            for t in range(T):
                want = host[q][t].clone()
                want[f:to] = src[f:to]
                assert torch.equal(dev[q][t].cpu(), want), (q, t)
        for dtype, op in [(torch.float32, N.SUM), (torch.int64, N.MAX)]:
            host, dev = _make(world, T, n, dtype, seed=12)
            plane.thread_segments("thread_reduce", dev, op, [(f, to, -1, -1, root[0], root[1])])
            torch.cuda.synchronize()
            plane.check()
            assert "relay" in plane.last(), plane.last()
            total = _total(host, op)
            for q in range(world):
                for t in range(T):
                    got = dev[q][t].cpu()
                    assert torch.equal(got[:f], host[q][t][:f]) and torch.equal(got[to:], host[q][t][to:]), (q, t)
                    if (q, t) == root:
                        assert torch.equal(got[f:to], total[f:to])
                    elif t != 0 or q == root[0]:
                        assert torch.equal(got, host[q][t]), (q, t)
    finally:
        plane.set_option("seg_relay_min_bytes", 4 << 20)
