"""Host arrays through the GPU plane (chunked H2D || fused kernel || D2H pipeline in the library):
results must be bit-identical to the CUDA-tensor path on the same inputs (TCS:1915, PCS:1733)."""
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


@pytest.mark.parametrize("world,T", [(1, 2), (1, 8), (2, 1), (2, 3), (2, 8), (4, 4)])
@pytest.mark.parametrize("chunk,n,f", [(64 << 10, 300001, 0), (64 << 10, 16384 * 5, 16384), (8 << 20, 70000, 3)])
def test_host_pipeline_matches_device_path(planes, world, T, chunk, n, f):
    plane = planes(world)
    plane.set_option("host_chunk_bytes", chunk)
    g = torch.Generator().manual_seed(n + world * 31 + T)
    host = [[torch.randn(n, generator=g, dtype=torch.float32).pin_memory() for _ in range(T)] for _ in range(world)]
    orig = [[h.clone() for h in row] for row in host]
    dev = [[h.cuda() for h in row] for row in host]
    # reference: the CUDA-tensor path on the same data
    if T == 1:
        plane.allreduce([row[0] for row in dev], N.SUM, f, n)
    else:
        plane.thread_allreduce(dev, N.SUM, f, n)
    torch.cuda.synchronize()
    plane.thread_allreduce_host(host, N.SUM, f, n)
    assert "host-pipeline" in plane.last(), plane.last()
    plane.check()
    for r in range(world):
        for t in range(T):
            assert torch.equal(host[r][t][f:], dev[r][t].cpu()[f:]), (r, t, plane.last())
            assert torch.equal(host[r][t][:f], orig[r][t][:f])


def test_host_pipeline_single_destination_and_int64(planes):
    plane = planes(2)
    plane.set_option("host_chunk_bytes", 32 << 10)
    T, n = 4, 50000
    host = [[torch.full((n,), (r * T + t + 1), dtype=torch.int64).pin_memory() for t in range(T)] for r in range(2)]
    plane.thread_allreduce_host(host, N.SUM, 0, n, fanout=False, dst=2)
    total = sum(range(1, 2 * T + 1))
    for r in range(2):
        assert torch.all(host[r][2] == total)
        for t in (0, 1, 3):
            assert torch.all(host[r][t] == r * T + t + 1)


def test_host_pipeline_pageable_memory_still_correct(planes):
    plane = planes(1)
    plane.set_option("host_chunk_bytes", 64 << 10)
    T, n = 3, 100000
    host = [[torch.full((n,), float(t + 1)) for t in range(T)]]
    plane.thread_allreduce_host(host, N.SUM, 0, n)
    for t in range(T):
        assert torch.all(host[0][t] == 6.0)
