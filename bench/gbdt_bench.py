#!/usr/bin/env python
"""K13 benchmark (BASELINE.json config 5: GBDT 1M rows x 2k features x 256 bins, histogram build
fused with reduceScatterArray). Rows are split evenly over the ranks. ours = ONE fused kernel
(shared-memory histograms flushed into the owner's shard over NVLink); baseline = atomic histogram
with torch.index_add_ into a local [F,256,2] buffer + NCCL reduce_scatter of it.
Roofline: max(bins bytes / HBM, shard bytes*(p-1)/p / 770 GB/s); the build is in practice bound by
shared-memory atomics, reported as updates/s. Launch under torchrun for N > 1."""
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
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--features", type=int, default=2048)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    lr = int(os.environ.get("LOCAL_RANK", 0))
    addr, port = os.environ.get("MASTER_ADDR", "127.0.0.1"), int(os.environ.get("MASTER_PORT", 29500))
    torch.cuda.set_device(lr)
    master = CommMaster(world, port + 37 if world > 1 else 0, quiet=True, work_dir=tempfile.gettempdir()).start() if rank == 0 else None
    comm = ProcessCommSlave("gbdt", addr, master.port if master else port + 37, rank_hint=rank, host="127.0.0.1")
    plane = comm.init_gpu(device=lr)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", lr))
    rows = (a.rows // world + 15) // 16 * 16
    F = a.features
    g = torch.Generator(device="cuda").manual_seed(rank)
    bins = torch.randint(0, 256, (F, rows), device="cuda", generator=g, dtype=torch.uint8)
    grad = torch.randn(rows, device="cuda", generator=g)
    hess = torch.rand(rows, device="cuda", generator=g)
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    fpr = (F + world - 1) // world

    def ours():
        return plane.gbdt_hist_reduce_scatter(bins, grad, hess)

    gh = torch.stack([grad, hess], 1).contiguous()  # [rows, 2]
    full = torch.zeros(fpr * world, 256, 2, device="cuda")
    shard = torch.empty(fpr, 256, 2, device="cuda")
    FC = 64  # features per index_add_ call (bounds the int64 index tensor)

    def baseline():
        full.zero_()
        flat = full.view(-1, 2)
        for f0 in range(0, F, FC):
            f1 = min(F, f0 + FC)
            idx = bins[f0:f1].long() + (torch.arange(f0, f1, device="cuda") * 256)[:, None]
            flat.index_add_(0, idx.view(-1), gh.repeat(f1 - f0, 1))
        if world > 1:
            dist.reduce_scatter_tensor(shard, full)
            return shard
        return full

    def timed(fn, iters):
        fn()
        torch.cuda.synchronize()
        comm.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return comm.allreduce(e0.elapsed_time(e1) / iters, dop, DMAX)

    s, _ = ours()
    b = baseline()
    torch.cuda.synchronize()
    n_own = max(0, min(F, (rank + 1) * fpr) - rank * fpr)
    ref = b[:n_own] if world > 1 else b[rank * fpr: rank * fpr + n_own]
    err = ((s[:n_own] - ref).abs().max() / (ref.abs().max() + 1e-9)).item()
    t_ours = timed(ours, a.iters)
    t_base = timed(baseline, max(1, a.iters // 2))
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm = peaks.get("hbm_gbs", 6650.0)
        t_roof = max(F * rows / (hbm * 1e9), fpr * 2048 * (world - 1) / 770e9) * 1e3
        print(json.dumps({"bench": "gbdt_hist_reduce_scatter", "n_gpus": world, "rows_total": rows * world,
                          "rows_per_rank": rows, "features": F, "ours_ms": round(t_ours, 3),
                          "index_add_nccl_ms": round(t_base, 3), "speedup": round(t_base / t_ours, 2),
                          "Gupdates_per_s_per_gpu": round(F * rows / t_ours / 1e6, 1),
                          "bins_GBps_per_gpu": round(F * rows / t_ours / 1e6, 1), "hbm_roofline_ms": round(t_roof, 4),
                          "frac_of_hbm_roofline": round(t_roof / t_ours, 3), "rel_err_vs_baseline": err,
                          "kernel": plane.last()}))
    if dist.is_initialized():
        dist.destroy_process_group()
    comm.close(0)
    if master:
        master.stop(timeout_s=10)
        master.close()


if __name__ == "__main__":
    main()
