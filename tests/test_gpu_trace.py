"""Per-collective statistics (MP4J_TRACE / plane.trace): CUDA-event timing per call, aggregated per
(collective, algorithm)."""
import pytest
import torch

from ytk_mp4j_b200 import _native as N

pytestmark = pytest.mark.gpu


def test_stats_per_collective_and_algorithm():
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    plane = EmuPlane(4, 0, heap_bytes=128 << 20, staging_bytes=4 << 20)
    try:
        assert plane.stats() == []  # tracing is off by default
        plane.trace(True)
        n = 1 << 20
        xs = [torch.ones(n, device="cuda") for _ in range(4)]
        for _ in range(3):
            plane.allreduce(xs, N.SUM, 0, n, algo=N.ALGO_TWOSHOT)
        small = [torch.ones(1000, device="cuda") for _ in range(4)]
        plane.allreduce(small, N.SUM, 0, 1000)
        plane.allgather(xs, [0, n // 4, n // 2, 3 * n // 4], [n // 4, n // 2, 3 * n // 4, n])
        st = {(r["collective"], r["algo"]): r for r in plane.stats()}
        two = st[("allreduce", "twoshot")]
        assert two["calls"] == 3 and two["bytes"] == 3 * n * 4
        assert 0 < two["min_us"] <= two["avg_us"] <= two["max_us"] and two["algbw_GBps"] > 0
        assert st[("allreduce", "oneshot")]["calls"] == 1
        assert any(k[0] == "allgather" for k in st)
        assert torch.all(xs[0] == 4.0 ** 3)
        plane.stats(reset=True)
        assert plane.stats() == []
        plane.trace(False)
        plane.allreduce(small, N.SUM, 0, 1000)
        assert plane.stats() == []
    finally:
        plane.close()
