"""CUDA-graph capture of the launch-bound paths: small one-shot allreduces, a two-shot allreduce and
the fused thread kernel are captured once and replayed with new inputs. Epoch counters live in device
memory and flags never reset, so a replay is an ordinary sequence of launches; the one-shot path keeps
its exit barrier under capture because the slot alternation is frozen into the graph."""
import pytest
import torch

from ytk_mp4j_b200 import _native as N

pytestmark = pytest.mark.gpu


def test_capture_and_replay_dense_and_thread():
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    W, T = 4, 2
    plane = EmuPlane(W, 0, heap_bytes=128 << 20, staging_bytes=4 << 20)
    try:
        small = [torch.zeros(1000, device="cuda") for _ in range(W)]
        big = plane.empty(1 << 18, N.F32)  # symmetric: zero-copy two-shot
        rows = [[torch.zeros(1 << 16, device="cuda") for _ in range(T)] for _ in range(W)]
        s = torch.cuda.Stream()
        # warm up outside capture (module loading, lazy allocations)
        with torch.cuda.stream(s):
            for _ in range(3):
                plane.allreduce(small, N.SUM, 0, 1000, stream=s)
            plane.allreduce(big, N.SUM, 0, 1 << 18, algo=N.ALGO_TWOSHOT, stream=s)
            plane.thread_allreduce(rows, N.SUM, 0, 1 << 16, stream=s)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            cs = torch.cuda.current_stream()
            for _ in range(3):  # odd count: consecutive replays start on the same one-shot slot
                plane.allreduce(small, N.SUM, 0, 1000, stream=cs)
            plane.allreduce(big, N.SUM, 0, 1 << 18, algo=N.ALGO_TWOSHOT, stream=cs)
            plane.thread_allreduce(rows, N.SUM, 0, 1 << 16, stream=cs)
        for rep in range(4):
            for r in range(W):
                small[r].fill_(float(r + 1 + rep))
                big[r].fill_(float(r + rep))
                for t in range(T):
                    rows[r][t].fill_(float(r * T + t + rep))
            g.replay()
            torch.cuda.synchronize()
            plane.check()
            s1 = sum(r + 1 + rep for r in range(W))
            want_small = float(s1 * W * W)  # three chained allreduces: x -> W*... each multiplies the sum by W
            assert torch.all(small[0] == want_small) and torch.all(small[W - 1] == want_small), (rep, small[0][:3])
            assert torch.all(big[1] == float(sum(r + rep for r in range(W))))
            want_rows = float(sum(r * T + t + rep for r in range(W) for t in range(T)))
            assert torch.all(rows[2][1] == want_rows) and torch.all(rows[0][0] == want_rows)
    finally:
        plane.close()
