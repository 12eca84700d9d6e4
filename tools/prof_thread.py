"""Short single-GPU driver for profiling the flagship kernel (fused thread allreduce).

Usage (under gpurun, see /opt/skills/guides/B200_PROFILING.md):
  python tools/prof_thread.py [--algo register|cluster] [--world 1] [--T 8] [--elems 67108864] [--iters 4]
world > 1 uses the emulated plane (all ranks' rows in one cooperative launch on one GPU)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from ytk_mp4j_b200 import _native as N  # noqa: E402
from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane, THREAD_ALGO  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--algo", default="auto")
ap.add_argument("--world", type=int, default=1)
ap.add_argument("--T", type=int, default=8)
ap.add_argument("--elems", type=int, default=64 << 20)
ap.add_argument("--iters", type=int, default=4)
a = ap.parse_args()

plane = EmuPlane(a.world, 0, heap_bytes=256 << 20, staging_bytes=8 << 20)
rows = [[torch.full((a.elems,), 1e-30, device="cuda") for _ in range(a.T)] for _ in range(a.world)]
torch.cuda.synchronize()
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(a.iters):
    ev0.record()
    plane.thread_allreduce(rows, N.SUM, 0, a.elems, thread_algo=THREAD_ALGO[a.algo])
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    traffic = 2 * a.T * a.elems * 4 * a.world
    print("iter %d: %.4f ms  %.1f GB/s HBM (algorithmic)  %s" % (i, ms, traffic / ms / 1e6, plane.last()))
plane.check()
plane.close()
