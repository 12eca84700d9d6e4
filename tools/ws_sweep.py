#!/usr/bin/env python
"""Sweeps the fused ThreadCommSlave allreduce kernels inside ONE communicator (no re-launch per point).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/ws_sweep.py --configs "ver:dbg:slot_kib:slots:nvls ..."

ver 1 = register formulation, 2 = TMA formulation; dbg = diagnostics bit mask (results are then wrong on
purpose: 1 no NVLink payload, 2 no cross-GPU flag waits, 4 no HBM streams, 8 phase B local only,
16 phase C does not read the slot, 32 phase A does not write the slot, 64 evict-first slot reads);
nvls: 0 = P2P, 1 = NVLS, -1 = auto. Device-timed, max over ranks.
"""
import argparse
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1:0:64:4:-1 2:0:64:4:-1")
    ap.add_argument("--elems", type=int, default=64 << 20)
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    import torch
    from ytk_mp4j_b200 import CommMaster, Operands, Operators, ThreadCommSlave
    from ytk_mp4j_b200 import _native as NV

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    port = int(os.environ.get("MASTER_PORT", "29500")) + 17
    torch.cuda.set_device(lr)
    master = None
    if rank == 0:
        master = CommMaster(world, port if world > 1 else 0, quiet=True, work_dir=tempfile.gettempdir()).start()
        port = master.port
    T, n = args.threads, args.elems
    comm = ThreadCommSlave("sweep", T, "127.0.0.1", port, rank_hint=rank, host="127.0.0.1")
    pcs = comm.processCommSlave
    plane = comm.init_gpu(device=lr)
    fop, SUM = Operands.FLOAT_OPERAND(), Operators.Float.SUM
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    bufs = [torch.full((n,), 1e-30 * (t + 1 + rank * T), dtype=torch.float32, device="cuda") for t in range(T)]
    rows = []
    for cfg in args.configs.split():
        f = [int(x) for x in cfg.split(":")]
        ver, dbg, kib, slots, nvls = f[:5]
        plane.set_option("ws_tail_split", f[5] if len(f) > 5 else 4)
        plane.set_option("ws_fa_gpu", f[6] if len(f) > 6 else 1)
        plane.set_option("ws_ver", ver)
        plane.set_option("ws_dbg", dbg)
        plane.set_option("thread_slot_bytes", kib << 10)
        plane.set_option("thread_slots", slots)
        algo = {-1: NV.ALGO_AUTO, 0: NV.ALGO_TWOSHOT, 1: NV.ALGO_NVLS}[nvls]
        ok = None
        if dbg & ~128 == 0:
            chk = [torch.full((1 << 20,), float(rank * T + t + 1), device="cuda") for t in range(T)]
            plane.thread_allreduce(chk, fop, SUM, 0, 1 << 20, algo=algo)
            torch.cuda.synchronize()
            P = world * T
            ok = all(bool(torch.all(c == P * (P + 1) / 2)) for c in chk)
        for _ in range(3):
            plane.thread_allreduce(bufs, fop, SUM, 0, n, algo=algo)
        torch.cuda.synchronize()
        pcs.barrier()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for s, e in ev:
            s.record()
            plane.thread_allreduce(bufs, fop, SUM, 0, n, algo=algo)
            e.record()
        torch.cuda.synchronize()
        ms = sorted(s.elapsed_time(e) for s, e in ev)
        med = pcs.allreduce(ms[len(ms) // 2], dop, DMAX)
        avg = pcs.allreduce(sum(ms) / len(ms), dop, DMAX)
        row = {"n_gpus": world, "cfg": cfg, "ms_median": round(med, 4), "ms_avg": round(avg, 4), "ok": ok,
               "kernel": plane.last()}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
        plane.check()
        if dbg & 128:
            from ws_emu_bench import print_trace
            tr = plane.ws_trace().cpu()
            pcs.barrier()
            for r in range(world):  # one rank after the other
                if r == rank:
                    print("--- stage trace of rank %d, block 0" % r, flush=True)
                    print_trace(tr)
                    sys.stdout.flush()
                pcs.barrier()
    pcs.barrier()
    comm.close(0)
    if master is not None:
        master.stop(timeout_s=10)
        master.close()
    if rank == 0 and args.out:
        with open(args.out, "a") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
