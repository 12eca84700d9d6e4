"""Launch-bound regime: per-call time of a small one-shot allreduce issued eagerly vs replayed from a
CUDA graph (emulated ranks on one GPU, so the number isolates launch overhead, not NVLink latency)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ytk_mp4j_b200 import _native as N  # noqa: E402
from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane  # noqa: E402

W, CALLS = int(os.environ.get("W", 2)), 32
plane = EmuPlane(W, 0, heap_bytes=64 << 20, staging_bytes=2 << 20)
for n in (256, 8192, 65536):
    xs = [torch.full((n,), 1e-30, device="cuda") for _ in range(W)]
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(CALLS):
            plane.allreduce(xs, N.SUM, 0, n, stream=s)
    s.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        e0.record(s)
        for _ in range(CALLS):
            plane.allreduce(xs, N.SUM, 0, n, stream=s)
        e1.record(s)
    s.synchronize()
    eager = e0.elapsed_time(e1) * 1e3 / CALLS
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        cs = torch.cuda.current_stream()
        for _ in range(CALLS):
            plane.allreduce(xs, N.SUM, 0, n, stream=cs)
    g.replay()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    graph = e0.elapsed_time(e1) * 1e3 / (5 * CALLS)
    print("allreduce fp32 n=%d (%d B), world=%d emulated: eager %.2f us/call, graph replay %.2f us/call  [%s]"
          % (n, n * 4, W, eager, graph, plane.last()))
plane.check()
plane.close()
