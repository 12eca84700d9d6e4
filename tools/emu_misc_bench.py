"""Kernel-cost numbers for the rows without a bandwidth table (K5, K6, K8) on emulated ranks of ONE GPU:
the "links" are local memory, so these isolate what the kernels themselves cost, not NVLink."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ytk_mp4j_b200 import _native as N  # noqa: E402
from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane  # noqa: E402

W, n = 4, 16 << 20  # 64 MiB fp32 per rank
plane = EmuPlane(W, 0, heap_bytes=768 << 20, staging_bytes=8 << 20)
xs = plane.empty(n, N.F32)
for x in xs:
    x.fill_(1e-3)
cnt = [n // W] * W
fr = [q * (n // W) for q in range(W)]
to = [f + c for f, c in zip(fr, cnt)]


def timed(name, fn, bytes_moved):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        fn()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 5
    print("%-28s %9.1f us  %7.1f GB/s per rank (algorithmic)  [%s]" % (name, us, bytes_moved / us / 1e3, plane.last()))


S = n * 4
timed("allreduce lossless two-shot", lambda: plane.allreduce(xs, N.SUM, 0, n, algo=N.ALGO_TWOSHOT), S)
timed("allreduce wire bf16", lambda: plane.allreduce_wire(xs, 0, n, N.WIRE_BF16), S)
timed("allreduce wire fp8", lambda: plane.allreduce_wire(xs, 0, n, N.WIRE_FP8), S)
timed("broadcast root=1", lambda: plane.broadcast(xs, 0, n, 1), S)
timed("gather root=2", lambda: plane.gather(xs, fr, to, 2), S)
timed("scatter root=2", lambda: plane.scatter(xs, fr, to, 2), S)
timed("reduce root=3", lambda: plane.reduce(xs, N.SUM, 0, n, 3), S)
plane.check()
plane.close()
