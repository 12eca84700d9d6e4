"""GPU numerics for the ThreadCommSlave kernels: fused thread-reduce + inter-GPU + fan-out
(register variant, all world sizes through emulation) and the cluster/DSMEM variant (intra-GPU),
against sequential PyTorch sums in the same (thread-id, then rank) order -> bit-exact for fp32."""
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


def expected(host, op=N.SUM):
    # host[rank][thread]; threads first (thread-id order), then ranks (rank order)
    per_rank = []
    for row in host:
        acc = row[0].clone()
        for x in row[1:]:
            acc = acc + x if op == N.SUM else torch.maximum(acc, x)
        per_rank.append(acc)
    acc = per_rank[0].clone()
    for x in per_rank[1:]:
        acc = acc + x if op == N.SUM else torch.maximum(acc, x)
    return acc


@pytest.mark.parametrize("world,T", [(1, 1), (1, 2), (1, 8), (2, 3), (4, 8), (8, 2)])
@pytest.mark.parametrize("n,f", [(4096, 0), (100003, 0), (262144 + 64, 64), (1001, 1)])
def test_thread_allreduce_f32(planes, world, T, n, f):
    plane = planes(world)
    g = torch.Generator().manual_seed(n + world * 131 + T)
    host = [[torch.randn(n, generator=g, dtype=torch.float32) for _ in range(T)] for _ in range(world)]
    dev = [[h.cuda() for h in row] for row in host]
    plane.thread_allreduce(dev, N.SUM, f, n)
    torch.cuda.synchronize()
    plane.check()
    want = expected([[h[f:] for h in row] for row in host])
    for r in range(world):
        for t in range(T):
            got = dev[r][t].cpu()
            assert torch.equal(got[f:], want), (r, t, plane.last())
            assert torch.equal(got[:f], host[r][t][:f])


@pytest.mark.parametrize("world", [1, 2, 4])
def test_thread_allreduce_multi_chunk_per_block(planes, world):
    # enough data that persistent blocks iterate over several chunks (slot reuse path)
    plane = planes(world)
    T, n = 4, 24 << 20 >> 2  # 24 MiB of fp32 per buffer
    host = [[torch.full((n,), float(r * T + t + 1)) for t in range(T)] for r in range(world)]
    dev = [[h.cuda() for h in row] for row in host]
    plane.thread_allreduce(dev, N.SUM, 0, n)
    torch.cuda.synchronize()
    plane.check()
    total = float(sum(range(1, world * T + 1)))
    for r in range(world):
        for t in range(T):
            assert torch.all(dev[r][t] == total), (r, t)


def test_thread_reduce_no_fanout_and_other_dtypes(planes):
    plane = planes(2)
    T, n = 3, 5000
    for dtype, code_op in ((torch.int32, N.BOR), (torch.float64, N.MAX), (torch.int8, N.SUM)):
        dev = [[torch.full((n,), (1 << t) if dtype != torch.float64 else float(t - r), dtype=dtype).cuda()
                for t in range(T)] for r in range(2)]
        before = [[x.clone() for x in row] for row in dev]
        plane.thread_allreduce(dev, code_op, 0, n, fanout=False, dst=1, local_only=True)
        torch.cuda.synchronize()
        for r in range(2):
            if code_op == N.BOR:
                assert torch.all(dev[r][1] == 7)
            elif code_op == N.MAX:
                assert torch.all(dev[r][1] == float(T - 1 - r))
            else:
                assert torch.all(dev[r][1] == 7)
            assert torch.equal(dev[r][0], before[r][0]) and torch.equal(dev[r][2], before[r][2])
    plane.check()


@pytest.mark.parametrize("T", [2, 4, 8, 16])
@pytest.mark.parametrize("n", [4096, (1 << 20) + 4 * 37, 1 << 22])
def test_thread_cluster_dsmem_variant(planes, T, n):
    plane = planes(1)
    g = torch.Generator().manual_seed(n + T)
    host = [[torch.randn(n, generator=g, dtype=torch.float32) for _ in range(T)]]
    dev = [[h.cuda() for h in host[0]]]
    plane.thread_allreduce(dev, N.SUM, 0, n, thread_algo=1)
    torch.cuda.synchronize()
    assert "cluster-dsmem" in plane.last()
    # cluster order: each CTA folds its own subset (t = cta, cta+CS, ...), then CTAs in rank order
    cs = 8 if T % 8 == 0 else (4 if T % 4 == 0 else 2)
    parts = []
    for c in range(cs):
        acc = host[0][c].clone()
        for j in range(1, T // cs):
            acc = acc + host[0][c + j * cs]
        parts.append(acc)
    want = parts[0].clone()
    for x in parts[1:]:
        want = want + x
    for t in range(T):
        assert torch.equal(dev[0][t].cpu(), want), t


@pytest.mark.parametrize("thread_algo,name", [(0, "local-256bit"), (3, "register")])
def test_thread_local_variants_agree(planes, thread_algo, name):
    plane = planes(1)
    T, n = 8, 1 << 20
    g = torch.Generator().manual_seed(5)
    host = [[torch.randn(n, generator=g, dtype=torch.float32) for _ in range(T)]]
    dev = [[h.cuda() for h in host[0]]]
    plane.thread_allreduce(dev, N.SUM, 0, n, thread_algo=thread_algo)
    torch.cuda.synchronize()
    assert name in plane.last()
    want = expected(host)
    for t in range(T):
        assert torch.equal(dev[0][t].cpu(), want)


@pytest.mark.parametrize("world,T", [(2, 8), (4, 4), (8, 2)])
@pytest.mark.parametrize("thread_algo,name", [(5, "fused-ws"), (6, "fused-pull"), (7, "fused-hybrid"), (4, "fused-tma"), (3, "fused-p2p")])
def test_thread_fused_variants_agree(planes, world, T, thread_algo, name):
    """Warp-specialised kernel vs chunk-synchronous fallback: identical bits, several chunks per block."""
    plane = planes(world)
    n = (6 << 20) + 64  # > grid * slot so blocks iterate and recycle their slot ring
    g = torch.Generator().manual_seed(world * 7 + T)
    host = [[torch.randn(n, generator=g, dtype=torch.float32) for _ in range(T)] for _ in range(world)]
    dev = [[h.cuda() for h in row] for row in host]
    for rep in range(2):  # second call exercises the persistent chunk counters
        plane.thread_allreduce(dev, N.SUM, 0, n, thread_algo=thread_algo)
        torch.cuda.synchronize()
        plane.check()
        assert name in plane.last(), plane.last()
        want = expected(host)
        for r in range(world):
            for t in range(T):
                assert torch.equal(dev[r][t].cpu(), want), (rep, r, t)
        host = [[want.clone() for _ in range(T)] for _ in range(world)]


@pytest.mark.parametrize("world,T", [(2, 1), (2, 3), (4, 5), (8, 8), (2, 64)])
@pytest.mark.parametrize("n,f", [(4096, 0), (100004, 0), ((3 << 20) + 8, 4), (1 << 20, 1 << 19)])
def test_thread_fused_tma_any_T(planes, world, T, n, f):
    """TMA formulation: run-time T, 16-byte alignment, partial last tiles / chunks, from > 0."""
    plane = planes(world)
    if T == 64:
        n = min(n, 1 << 18)
        f = min(f, n // 2)
    g = torch.Generator().manual_seed(n + world * 17 + T)
    host = [[torch.randn(n, generator=g, dtype=torch.float32) for _ in range(T)] for _ in range(world)]
    dev = [[h.cuda() for h in row] for row in host]
    plane.thread_allreduce(dev, N.SUM, f, n, thread_algo=4)
    torch.cuda.synchronize()
    plane.check()
    assert "fused-tma" in plane.last(), plane.last()
    want = expected([[h[f:] for h in row] for row in host])
    for r in range(world):
        for t in range(T):
            got = dev[r][t].cpu()
            assert torch.equal(got[f:], want), (r, t)
            assert torch.equal(got[:f], host[r][t][:f])


def test_thread_fused_tma_no_fanout_f64_max(planes):
    plane = planes(2)
    T, n = 3, 70000
    g = torch.Generator().manual_seed(3)
    host = [[torch.randn(n, generator=g, dtype=torch.float64) for _ in range(T)] for _ in range(2)]
    dev = [[h.cuda() for h in row] for row in host]
    plane.thread_allreduce(dev, N.MAX, 0, n, fanout=False, dst=2, thread_algo=4)
    torch.cuda.synchronize()
    plane.check()
    want = expected(host, N.MAX)
    for r in range(2):
        assert torch.equal(dev[r][2].cpu(), want)
        assert torch.equal(dev[r][0].cpu(), host[r][0]) and torch.equal(dev[r][1].cpu(), host[r][1])
