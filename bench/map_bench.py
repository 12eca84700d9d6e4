#!/usr/bin/env python
"""Sparse-map benchmark (BASELINE.json metric "sparse map keys/s", config 4: 10M hashed features).

Every rank holds n (key, value) pairs drawn from a space of --features hashed feature ids (so ranks
overlap heavily, like minibatch gradients of a sparse LR) and calls allreduce_map. Reported:
keys/s = total input pairs over all ranks / time, for ours (fused partition -> peer push -> hash
merge -> allgather push) and for a baseline built from library pieces on the same GPUs:
torch.unique + index_add_ locally, NCCL all_gather of the padded pair lists, torch.unique +
index_add_ again (what a user would write without a map collective).
Launch under torchrun for N > 1."""
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
    ap.add_argument("--features", type=int, default=10_000_000)
    ap.add_argument("--n", type=int, default=2_000_000)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    lr = int(os.environ.get("LOCAL_RANK", 0))
    addr, port = os.environ.get("MASTER_ADDR", "127.0.0.1"), int(os.environ.get("MASTER_PORT", 29500))
    torch.cuda.set_device(lr)
    master = CommMaster(world, port + 31 if world > 1 else 0, quiet=True, work_dir=tempfile.gettempdir()).start() if rank == 0 else None
    comm = ProcessCommSlave("map", addr, master.port if master else port + 31, rank_hint=rank, host="127.0.0.1")
    plane = comm.init_gpu(device=lr, heap_bytes=4 << 30)
    if world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", lr))
    g = torch.Generator(device="cuda").manual_seed(rank + 1)
    keys = torch.randperm(a.features, device="cuda", generator=g)[: a.n].to(torch.int64)
    vals = torch.randn(a.n, device="cuda", generator=g)
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    out_cap = min(a.features, a.n * world)
    max_n = a.n

    def ours():
        return plane._map_collective("allreduce", [keys], [vals], max_n, out_cap, Operators.Float.SUM.code, "mix", None,
                                     clone=False)[0]

    def baseline():
        if world > 1:
            ak = torch.empty(world * a.n, dtype=torch.int64, device="cuda")
            av = torch.empty(world * a.n, device="cuda")
            dist.all_gather_into_tensor(ak, keys)
            dist.all_gather_into_tensor(av, vals)
        else:
            ak, av = keys, vals
        uk, inv = torch.unique(ak, return_inverse=True)
        return uk, torch.zeros(uk.numel(), device="cuda").index_add_(0, inv, av)

    def timed(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        comm.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return comm.allreduce(e0.elapsed_time(e1) / a.iters, dop, DMAX)

    k, v, _ = ours()
    bk, bv = baseline()
    ok = k.numel() == bk.numel()
    order = torch.argsort(k)
    err = (v[order] - bv).abs().max().item() if ok else float("nan")
    t_ours, t_base = timed(ours), timed(baseline)
    if rank == 0:
        total = a.n * world
        print(json.dumps({"bench": "allreduce_map", "n_gpus": world, "pairs_per_rank": a.n, "feature_space": a.features,
                          "unique_keys": int(k.numel()), "ours_ms": round(t_ours, 3), "baseline_ms": round(t_base, 3),
                          "ours_Mkeys_per_s": round(total / t_ours / 1e3, 1),
                          "baseline_Mkeys_per_s": round(total / t_base / 1e3, 1), "speedup": round(t_base / t_ours, 2),
                          "same_key_set": ok, "max_abs_err_vs_baseline": err, "kernel": plane.last()}))
    if dist.is_initialized():
        dist.destroy_process_group()
    comm.close(0)
    if master:
        master.stop(timeout_s=10)
        master.close()


if __name__ == "__main__":
    main()
