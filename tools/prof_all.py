"""Single-GPU driver that launches every kernel family a few times, for `ncu -k regex:<kernel>` captures
(tools/prof_all.sh) and for a plain timing list. Multi-rank kernels run on EMULATED ranks (one cooperative launch,
all ranks' blocks on this GPU), as the profiling guide requires for kernels that wait on flags.

  python tools/prof_all.py --which grs_twoshot|grs_oneshot|ws|seg_rs|seg_ag|map|gbdt|codec|lr|lr_residual|all
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from ytk_mp4j_b200 import _native as N  # noqa: E402
from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane, lr_softmax_residual  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--which", default="all")
ap.add_argument("--world", type=int, default=2)
ap.add_argument("--iters", type=int, default=3)
a = ap.parse_args()
W = a.world
plane = EmuPlane(W, 0, heap_bytes=1 << 30, staging_bytes=8 << 20)
g = torch.Generator(device="cuda").manual_seed(0)


def timed(name, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(a.iters):
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        print("%-14s iter %d: %.4f ms  %s" % (name, i, e0.elapsed_time(e1), plane.last()))
    plane.check()


def want(k):
    return a.which in ("all", k)


if want("grs_twoshot"):
    n = 64 << 20  # 256 MiB fp32 per rank, symmetric (zero copy)
    xs = plane.empty(n, N.F32)
    for x in xs:
        x.fill_(1e-30)
    timed("grs_twoshot", lambda: plane.allreduce(xs, N.SUM, 0, n, algo=N.ALGO_TWOSHOT))
if want("grs_oneshot"):
    xs1 = [torch.full((16384,), 1.0, device="cuda") for _ in range(W)]
    timed("grs_oneshot", lambda: plane.allreduce(xs1, N.SUM, 0, 16384, algo=N.ALGO_ONESHOT))
if want("ws"):
    n = (64 << 20) // W
    rows = [[torch.full((n,), 1e-30, device="cuda") for _ in range(8)] for _ in range(W)]
    timed("ws", lambda: plane.thread_allreduce(rows, N.SUM, 0, n))
    del rows
if want("seg_rs") or want("seg_ag"):
    n = (64 << 20) // W
    rows = [[torch.full((n,), 1e-30, device="cuda") for _ in range(8)] for _ in range(W)]
    w = n // (W * 8)
    sl = [(q, t, (q * 8 + t) * w, (q * 8 + t + 1) * w) for q in range(W) for t in range(8)]
    if want("seg_rs"):
        timed("seg_rs", lambda: plane.thread_segments("thread_reduce_scatter", rows, N.SUM,
                                                      [(lo, hi, -1, -1, q, t) for q, t, lo, hi in sl]))
    if want("seg_ag"):
        timed("seg_ag", lambda: plane.thread_segments("thread_allgather", rows, N.COPY,
                                                      [(lo, hi, q, t, -1, -1) for q, t, lo, hi in sl]))
    del rows
if want("map"):
    keys = [torch.randint(0, 10_000_000, (2_000_000,), device="cuda", generator=g) for _ in range(W)]
    vals = [torch.ones(2_000_000, device="cuda") for _ in range(W)]
    timed("map", lambda: plane.map_collective("allreduce", keys, vals, N.SUM))
if want("gbdt"):
    F, rows_g = 2048, 1_000_000 // W
    bins = [torch.randint(0, 256, (F, rows_g), device="cuda", generator=g, dtype=torch.uint8) for _ in range(W)]
    gr = [torch.randn(rows_g, device="cuda", generator=g) for _ in range(W)]
    he = [torch.rand(rows_g, device="cuda", generator=g) + 0.1 for _ in range(W)]
    timed("gbdt", lambda: plane.gbdt_hist_reduce_scatter(bins, gr, he))
    del bins
if want("codec"):
    xs2 = plane.empty(16 << 20, N.F32)
    for x in xs2:
        x.fill_(0.5)
    timed("codec_bf16", lambda: plane.allreduce_wire(xs2, 0, 16 << 20, N.WIRE_BF16))
if want("lr"):
    F, Cn, K = 8192, 128, 65536
    xts = [torch.randn(F, K, device="cuda", generator=g).to(torch.bfloat16) for _ in range(1)] * W
    rts = [torch.randn(Cn, K, device="cuda", generator=g).to(torch.bfloat16) for _ in range(1)] * W
    timed("lr_xtr", lambda: plane.lr_xtr_reduce_scatter(xts, rts))
if want("lr_residual"):
    logits = torch.randn(65536, 128, device="cuda", generator=g).to(torch.bfloat16)
    y = torch.randint(0, 128, (65536,), device="cuda", generator=g)
    rt = torch.empty(128, 65536, dtype=torch.bfloat16, device="cuda")
    loss = torch.zeros(1, device="cuda")
    timed("lr_residual", lambda: lr_softmax_residual(logits, y, rt, loss))
plane.close()
