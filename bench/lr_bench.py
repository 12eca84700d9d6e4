#!/usr/bin/env python
"""K12 benchmark: fused tcgen05 X^T R GEMM + reduce-scatter vs cuBLAS GEMM + NCCL reduce_scatter
(BASELINE.json config 4, dense part). Launch under torchrun for N > 1, plain python for N = 1.
Roofline: max(2*F*C*K / measured GEMM peak, Xt bytes / measured HBM, shard bytes*(p-1)/p / 770 GB/s)."""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from ytk_mp4j_b200 import CommMaster, Operands, Operators, ProcessCommSlave  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--F", type=int, default=8192)
    ap.add_argument("--C", type=int, default=128)
    ap.add_argument("--K", type=int, default=65536)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    lr = int(os.environ.get("LOCAL_RANK", 0))
    addr, port = os.environ.get("MASTER_ADDR", "127.0.0.1"), int(os.environ.get("MASTER_PORT", 29500))
    torch.cuda.set_device(lr)
    master = CommMaster(world, port + 29 if world > 1 else 0, quiet=True, work_dir=tempfile.gettempdir()).start() if rank == 0 else None
    mport = master.port if master else port + 29
    comm = ProcessCommSlave("lr", addr, mport, rank_hint=rank, host="127.0.0.1")
    plane = comm.init_gpu(device=lr)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", lr))
    g = torch.Generator(device="cuda").manual_seed(rank)
    # several Xt copies so consecutive iterations do not find their operand in L2 (126 MB)
    ncopy = max(1, min(4, (600 << 20) // (a.F * a.K * 2) + 1))
    xts = [torch.randn(a.F, a.K, device="cuda", generator=g).to(torch.bfloat16) for _ in range(ncopy)]
    rt = torch.randn(a.C, a.K, device="cuda", generator=g).to(torch.bfloat16)
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX

    def timed(fn):
        for i in range(3):
            fn(i)
        torch.cuda.synchronize()
        comm.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.iters):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        return comm.allreduce(e0.elapsed_time(e1) / a.iters, dop, DMAX)

    def ours(i):
        plane.lr_xtr_reduce_scatter(xts[i % ncopy], rt)

    fpr = (a.F + world - 1) // world
    outbuf = torch.empty(fpr * world, a.C, device="cuda")
    shard = torch.empty(fpr, a.C, device="cuda")

    def baseline(i):
        torch.matmul(xts[i % ncopy][: a.F].float() if False else xts[i % ncopy], rt.t(), out=None)
        gfull = torch.matmul(xts[i % ncopy], rt.t()).float()
        if world > 1:
            outbuf[: a.F] = gfull
            dist.reduce_scatter_tensor(shard, outbuf)

    def baseline2(i):
        gfull = (xts[i % ncopy] @ rt.t()).float()  # cuBLAS bf16 GEMM, fp32 copy for the collective
        if world > 1:
            outbuf[: a.F] = gfull
            dist.reduce_scatter_tensor(shard, outbuf)

    # correctness spot check against the baseline
    s, _ = plane.lr_xtr_reduce_scatter(xts[0], rt)
    baseline2(0)
    torch.cuda.synchronize()
    ref = shard if world > 1 else (xts[0] @ rt.t()).float()[:fpr]
    err = (s[: ref.shape[0]] - ref).abs().max().item() / (ref.abs().max().item() + 1e-9)
    t_ours = timed(ours)
    t_base = timed(baseline2)
    if rank == 0:
        flops = 2.0 * a.F * a.C * a.K
        bytes_x = a.F * a.K * 2 + a.C * a.K * 2
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm, tf = peaks.get("hbm_gbs", 6650.0), peaks.get("bf16_tflops", 1590.0)
        t_roof = max(flops / (tf * 1e12), bytes_x / (hbm * 1e9),
                     fpr * a.C * 4 * (world - 1) / 770e9) * 1e3
        print(json.dumps({"bench": "lr_xtr_reduce_scatter", "n_gpus": world, "F": a.F, "C": a.C, "K": a.K,
                          "ours_ms": round(t_ours, 4), "cublas_nccl_ms": round(t_base, 4),
                          "speedup": round(t_base / t_ours, 3), "tflops": round(flops / t_ours / 1e9, 1),
                          "xt_GBps": round(bytes_x / t_ours / 1e6, 1), "roofline_ms": round(t_roof, 4),
                          "frac_of_roofline": round(t_roof / t_ours, 3), "rel_err_vs_baseline": err,
                          "l2_policy": "%d operand copies rotate (> L2)" % ncopy, "kernel": plane.last()}))
    if dist.is_initialized():
        dist.destroy_process_group()
    comm.close(0)
    if master:
        master.stop(timeout_s=10)
        master.close()


if __name__ == "__main__":
    main()
