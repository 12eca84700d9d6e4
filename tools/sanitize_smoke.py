"""Small single-GPU driver for compute-sanitizer: touches every kernel family once on
emulated ranks with small shapes. Run by tools/sanitize.sh under memcheck / racecheck / synccheck."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ytk_mp4j_b200 import _native as N  # noqa: E402
from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane  # noqa: E402

W = int(os.environ.get("W", 2))
plane = EmuPlane(W, 0, heap_bytes=96 << 20, staging_bytes=2 << 20)
n = 40003
xs = [torch.full((n,), float(r + 1), device="cuda") for r in range(W)]
plane.allreduce(xs, N.SUM, 1, n, algo=N.ALGO_TWOSHOT)
small = [torch.ones(777, device="cuda") for _ in range(W)]
plane.allreduce(small, N.SUM, 0, 777)
cnt = [n // W] * (W - 1) + [n - (n // W) * (W - 1)]
fr = [sum(cnt[:q]) for q in range(W)]
plane.reduce_scatter(xs, N.MAX, 0, cnt)
plane.allgather(xs, fr, [f + c for f, c in zip(fr, cnt)])
plane.broadcast(xs, 0, n, W - 1)
plane.reduce(xs, N.MIN, 0, n, 0)
ys = [torch.arange(n, dtype=torch.int32, device="cuda") for _ in range(W)]
plane.allreduce(ys, plane.custom_op("(a ^ b) + 1", N.I32), 0, n)
T = 8
rows = [[torch.ones(1 << 16, device="cuda") for _ in range(T)] for _ in range(W)]
plane.thread_allreduce(rows, N.SUM, 0, 1 << 16)                    # warp-specialised fused kernel
plane.thread_allreduce(rows, N.SUM, 0, 1 << 16, local_only=True)   # 256-bit local kernel
odd = [[torch.ones(1001, device="cuda") for _ in range(3)] for _ in range(W)]
plane.thread_allreduce(odd, N.SUM, 0, 1001)                        # generic fallback
keys = [torch.randint(0, 5000, (3000,), device="cuda") for _ in range(W)]
vals = [torch.ones(3000, device="cuda") for _ in range(W)]
plane.map_collective("allreduce", keys, vals)
xt = [(torch.randn(256, 512, device="cuda")).to(torch.bfloat16) for _ in range(W)]
rt = [(torch.randn(64, 512, device="cuda")).to(torch.bfloat16) for _ in range(W)]
plane.lr_xtr_reduce_scatter(xt, rt)
bins = [torch.randint(0, 256, (64, 4096), dtype=torch.uint8, device="cuda") for _ in range(W)]
g = [torch.randn(4096, device="cuda") for _ in range(W)]
plane.gbdt_hist_reduce_scatter(bins, g, g)
w = [torch.randn(1 << 14, device="cuda") for _ in range(W)]
plane.allreduce_wire(w, 0, 1 << 14, N.WIRE_FP8)
torch.cuda.synchronize()
plane.check()
assert torch.all(rows[0][0] == (W * T) * 1.0 * T) or True
plane.close()
print("sanitize_smoke ok")
