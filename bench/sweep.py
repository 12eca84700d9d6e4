#!/usr/bin/env python
"""Dense collective sweep: ours (every algorithm) vs NCCL on the same GPUs (BASELINE.json config 2:
"dense float allreduce/reduceScatter/allgather sweep 1 KB-1 GB at 2/4/8 B200 vs ... NCCL").

Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
            --master-port P bench/sweep.py [--max-bytes 1073741824] [--out gpurun_out/sweep_N.jsonl]
Timing: CUDA events on the launching stream, `iters` back-to-back calls after warm-up, max over
ranks (through our own host allreduce). Buffers rotate through a 512 MiB symmetric pool so that
consecutive iterations of sizes below the L2 size do not hit the same lines.
busbw convention (nccl-tests): allreduce algbw*2(p-1)/p; reduce-scatter / allgather algbw*(p-1)/p.
"""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from ytk_mp4j_b200 import CommMaster, Operands, Operators, ProcessCommSlave  # noqa: E402
from ytk_mp4j_b200 import _native as N  # noqa: E402

ALGOS = {"auto": N.ALGO_AUTO, "oneshot": N.ALGO_ONESHOT, "twoshot": N.ALGO_TWOSHOT, "nvls": N.ALGO_NVLS}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", type=int, default=1024)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--factor", type=int, default=4)
    ap.add_argument("--out", default=None)
    ap.add_argument("--dtype", default="float", choices=["float", "double"])
    ap.add_argument("--no-nccl", action="store_true")
    ap.add_argument("--custom", action="store_true",
                    help="also time allreduce with the NVRTC-compiled custom operator `a + b` (two-shot P2P)")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    lr = int(os.environ.get("LOCAL_RANK", 0))
    addr, port = os.environ.get("MASTER_ADDR", "127.0.0.1"), int(os.environ.get("MASTER_PORT", 29500))
    torch.cuda.set_device(lr)
    master = CommMaster(world, port + 23, quiet=True, work_dir=tempfile.gettempdir()).start() if rank == 0 else None
    comm = ProcessCommSlave("sweep", addr, port + 23, rank_hint=rank, host="127.0.0.1")
    pool_bytes = max(512 << 20, 2 * a.max_bytes)
    plane = comm.init_gpu(device=lr, heap_bytes=pool_bytes + (512 << 20))
    if not a.no_nccl and world > 1:
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", lr))
    operand = Operands.FLOAT_OPERAND() if a.dtype == "float" else Operands.DOUBLE_OPERAND()
    tdt = torch.float32 if a.dtype == "float" else torch.float64
    es = 4 if a.dtype == "float" else 8
    SUM = Operators.Float.SUM if a.dtype == "float" else Operators.Double.SUM
    pool = comm.empty(pool_bytes // es, operand)
    pool.fill_(1e-30)
    npool = torch.full((pool_bytes // es,), 1e-30, dtype=tdt, device="cuda")
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    rows = []
    algos = ["auto", "oneshot", "twoshot"] + (["nvls"] if plane.has_multicast else [])

    def timed(fn, nbytes):
        iters = 20 if nbytes <= (64 << 20) else 8
        warm = 5 if nbytes <= (64 << 20) else 2
        nreg = max(1, min(64, (256 << 20) // max(nbytes, 1)))
        for i in range(warm):
            fn(i % nreg)
        torch.cuda.synchronize()
        comm.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            fn(i % nreg)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / iters
        return comm.allreduce(us, dop, DMAX)

    size = a.min_bytes
    while size <= a.max_bytes:
        n = size // es
        counts = [n // world] * (world - 1) + [n - (n // world) * (world - 1)]
        froms = [sum(counts[:q]) for q in range(world)]
        tos = [f + c for f, c in zip(froms, counts)]
        stride = ((n + 63) // 64) * 64
        for coll in ("allreduce", "reduce_scatter", "allgather"):
            factor = 2 * (world - 1) / world if coll == "allreduce" else (world - 1) / world
            for algo in algos:
                if coll != "allreduce" and algo == "oneshot":
                    continue
                if algo == "oneshot" and size > (256 << 10):
                    continue

                def ours(i, coll=coll, algo=algo):
                    x = pool[i * stride:i * stride + n]
                    if coll == "allreduce":
                        plane.allreduce(x, operand, SUM, 0, n, algo=ALGOS[algo])
                    elif coll == "reduce_scatter":
                        plane.reduce_scatter(x, operand, SUM, 0, counts, algo=ALGOS[algo])
                    else:
                        plane.allgather(x, operand, froms, tos, algo=ALGOS[algo])
                us = timed(ours, size)
                rows.append({"impl": "ours", "coll": coll, "algo": algo, "bytes": size, "us": round(us, 2),
                             "algbw_GBps": round(size / us / 1e3, 2), "busbw_GBps": round(size / us / 1e3 * factor, 2),
                             "picked": plane.last()})
            if a.custom and coll == "allreduce":
                cop = Operators.device_expr("a + b")

                def ours_custom(i):
                    x = pool[i * stride:i * stride + n]
                    plane.allreduce(x, operand, cop, 0, n, algo=N.ALGO_TWOSHOT if size > (256 << 10) else N.ALGO_AUTO)
                us = timed(ours_custom, size)
                rows.append({"impl": "ours", "coll": coll, "algo": "custom", "bytes": size, "us": round(us, 2),
                             "algbw_GBps": round(size / us / 1e3, 2), "busbw_GBps": round(size / us / 1e3 * factor, 2),
                             "picked": plane.last()})
            if not a.no_nccl and world > 1:
                m = n // world

                def nccl(i, coll=coll):
                    x = npool[i * stride:i * stride + n]
                    if coll == "allreduce":
                        dist.all_reduce(x)
                    elif coll == "reduce_scatter":
                        dist.reduce_scatter_tensor(x[rank * m:(rank + 1) * m], x[:m * world])
                    else:
                        dist.all_gather_into_tensor(x[:m * world], x[rank * m:(rank + 1) * m])
                us = timed(nccl, size)
                rows.append({"impl": "nccl", "coll": coll, "algo": "nccl", "bytes": size, "us": round(us, 2),
                             "algbw_GBps": round(size / us / 1e3, 2), "busbw_GBps": round(size / us / 1e3 * factor, 2)})
        size *= a.factor
    plane.check()
    comm.barrier()
    if rank == 0:
        out = a.out or "gpurun_out/sweep_%d.jsonl" % world
        os.makedirs(os.path.dirname(out) or ".", exist_ok=True)
        with open(out, "w") as f:
            for r in rows:
                r["n_gpus"] = world
                f.write(json.dumps(r) + "\n")
        # compact table on stdout
        print("%-15s %-8s %12s %10s %10s" % ("coll", "algo", "bytes", "us", "busbw"))
        for r in rows:
            print("%-15s %-8s %12d %10.1f %10.1f" % (r["coll"], r["impl"][0] + ":" + r["algo"], r["bytes"], r["us"],
                                                      r["busbw_GBps"]))
    if dist.is_initialized():
        dist.destroy_process_group()
    comm.close(0)
    if master is not None:
        master.stop(timeout_s=10)
        master.close()


if __name__ == "__main__":
    main()
