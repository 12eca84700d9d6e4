"""Short single-GPU driver for profiling the K12 kernel (tcgen05 GEMM + red.add epilogue)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane  # noqa: E402

F, C, K = 8192, int(os.environ.get("C", 128)), 65536
plane = EmuPlane(1, 0, heap_bytes=256 << 20, staging_bytes=4 << 20)
xt = torch.randn(F, K, device="cuda").to(torch.bfloat16)
rt = torch.randn(C, K, device="cuda").to(torch.bfloat16)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(4):
    e0.record()
    plane.lr_xtr_reduce_scatter([xt], [rt])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print("iter %d: %.4f ms  %.1f TFLOP/s  %.1f GB/s Xt" % (i, ms, 2.0 * F * C * K / ms / 1e9, (F + C) * K * 2 / ms / 1e6))
plane.close()
