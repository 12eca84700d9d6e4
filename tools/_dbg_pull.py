import sys, os
sys.path.insert(0, os.getcwd())
import torch
from ytk_mp4j_b200 import _native as N
from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
W, T = 2, 8
plane = EmuPlane(W, 0, heap_bytes=192 << 20, staging_bytes=8 << 20)
for algo in (5, 6, 6, 6):
    for n in (16384 * 10, 16384 * 74, 16384 * 74 * 2, 16384 * 74 * 4, 16384 * 74 * 5, (6 << 20) + 64):
        g = torch.Generator().manual_seed(n)
        host = [[torch.randint(0, 1 << 20, (n,), generator=g).float() for t in range(T)] for r in range(W)]
        dev = [[h.cuda() for h in row] for row in host]
        plane.thread_allreduce(dev, N.SUM, 0, n, thread_algo=algo)
        torch.cuda.synchronize()
        plane.check()
        want = sum(sum(row) for row in host)
        for r in range(W):
            for t in (0, T - 1):
                got = dev[r][t].cpu()
                bad = (got != want).nonzero().flatten()
                if bad.numel():
                    uniq = []
                    chunks = torch.unique(bad // 16384)
                    print("algo", algo, "n", n, "rank", r, "t", t, "bad", bad.numel(), "first", int(bad[0]), "last", int(bad[-1]),
                          "nchunks_bad", chunks.numel(), "chunks", chunks[:12].tolist(),
                          "in-chunk off", int(bad[0]) % 16384, int(bad[-1]) % 16384)
        print("algo", algo, "n", n, "done", plane.last())
