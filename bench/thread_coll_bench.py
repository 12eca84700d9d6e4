#!/usr/bin/env python
"""Device time of the seven ThreadCommSlave array collectives on CUDA tensors: T buffers of n fp32 per GPU
(default 8 x 64 Mi, BASELINE.json config 3), one fused launch each. Run alone (1 GPU) or under torchrun.

  python bench/thread_coll_bench.py [--elems N] [--threads T] [--steps K]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 bench/thread_coll_bench.py
"""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--elems", type=int, default=64 << 20)
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--out", default="")
    ap.add_argument("--only", default="", help="comma-separated subset of the collectives")
    args = ap.parse_args()
    import torch
    from ytk_mp4j_b200 import CommMaster, Operands, Operators, ThreadCommSlave

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    port = int(os.environ.get("MASTER_PORT", "29500")) + 19
    torch.cuda.set_device(lr)
    master = None
    if rank == 0:
        master = CommMaster(world, port if world > 1 else 0, quiet=True, work_dir=tempfile.gettempdir()).start()
        port = master.port
    T, n = args.threads, args.elems
    comm = ThreadCommSlave("tcb", T, "127.0.0.1", port, rank_hint=rank, host="127.0.0.1")
    pcs = comm.processCommSlave
    plane = comm.init_gpu(device=lr)
    fop, SUM = Operands.FLOAT_OPERAND(), Operators.Float.SUM
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    bufs = [torch.full((n,), 1e-30 * (t + 1 + rank * T), dtype=torch.float32, device="cuda") for t in range(T)]
    P = world * T
    w = n // P
    slices = [(q, t, (q * T + t) * w, (q * T + t + 1) * w) for q in range(world) for t in range(T)]
    cases = {
        "allreduce": lambda: plane.thread_allreduce(bufs, fop, SUM, 0, n),
        "reduce_scatter": lambda: plane.thread_segments("thread_reduce_scatter", bufs, fop, SUM,
                                                        [(lo, hi, -1, -1, q, t) for q, t, lo, hi in slices]),
        "reduce": lambda: plane.thread_segments("thread_reduce", bufs, fop, SUM, [(0, n, -1, -1, world - 1, T - 1)]),
        "allgather": lambda: plane.thread_segments("thread_allgather", bufs, fop, None,
                                                   [(lo, hi, q, t, -1, -1) for q, t, lo, hi in slices]),
        "gather": lambda: plane.thread_segments("thread_gather", bufs, fop, None,
                                                [(lo, hi, q, t, 0, 0) for q, t, lo, hi in slices]),
        "scatter": lambda: plane.thread_segments("thread_scatter", bufs, fop, None,
                                                 [(lo, hi, 0, 0, q, t) for q, t, lo, hi in slices]),
        "broadcast": lambda: plane.thread_segments("thread_broadcast", bufs, fop, None, [(0, n, 0, 0, -1, -1)]),
    }
    # algorithmic HBM bytes per GPU (reads + writes of the thread buffers)
    S = n * 4
    hbm = {"allreduce": 2 * T * S, "reduce_scatter": T * S + S / world, "reduce": T * S + S / world,
           "allgather": S / world + T * S, "gather": S / world + S / world, "scatter": S / world + S / world,
           "broadcast": S / world + T * S}
    rows = []
    only = [x for x in args.only.split(",") if x]
    for name, fn in cases.items():
        if only and name not in only:
            continue
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        pcs.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for s, e in ev:
            s.record()
            fn()
            e.record()
        torch.cuda.synchronize()
        plane.check()
        ms = sorted(s.elapsed_time(e) for s, e in ev)
        med = pcs.allreduce(ms[len(ms) // 2], dop, DMAX)
        row = {"n_gpus": world, "T": T, "elems": n, "collective": name, "ms_median": round(med, 4),
               "algorithmic_hbm_GBps_per_gpu": round(hbm[name] / (med * 1e-3) / 1e9, 1), "kernel": plane.last()}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    pcs.barrier()
    comm.close(0)
    if master is not None:
        master.stop(timeout_s=10)
        master.close()
    if rank == 0 and args.out:
        with open(args.out, "a") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
