#!/usr/bin/env python
"""Fused thread-allreduce kernels on EMULATED ranks of one GPU (no NVLink latency): isolates what the
kernel structure itself costs. Total HBM traffic equals the 1-GPU flagship (world x T x n x 4 B = 2 GiB in)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


STAGES = ["A done", "sig saw A", "FA released", "B saw FA", "B data done", "FB released", "C saw FB", "C slot read"]


def print_trace(tr, first=4, count=12):
    """Per-chunk stage times of block 0, microseconds relative to the chunk's 'A done'."""
    t0 = int(tr[tr > 0].min()) if bool((tr > 0).any()) else 0
    print("chunk  A_done(abs us) | " + " | ".join(STAGES[1:]) + "   (us after A done)")
    for k in range(first, first + count):
        row = tr[k].tolist()
        if row[0] == 0:
            continue
        print("%5d  %10.1f | " % (k, (row[0] - t0) / 1e3) + " | ".join("%8.1f" % ((x - row[0]) / 1e3) for x in row[1:]))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--configs", default="5:0:64:4 4:0:64:4")
    ap.add_argument("--steps", type=int, default=10)
    args = ap.parse_args()
    import torch
    from ytk_mp4j_b200 import _native as N
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    W, T = args.world, args.threads
    n = (64 << 20) // W
    plane = EmuPlane(W, 0, heap_bytes=512 << 20, staging_bytes=8 << 20)
    dev = [[torch.full((n,), 1e-30, dtype=torch.float32, device="cuda") for _ in range(T)] for _ in range(W)]
    for cfg in args.configs.split():
        talgo, dbg, kib, slots = (int(x) for x in cfg.split(":"))
        plane.set_option("ws_dbg", dbg)
        plane.set_option("thread_slot_bytes", kib << 10)
        plane.set_option("thread_slots", slots)
        for _ in range(3):
            plane.thread_allreduce(dev, N.SUM, 0, n, thread_algo=talgo)
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for s, e in ev:
            s.record()
            plane.thread_allreduce(dev, N.SUM, 0, n, thread_algo=talgo)
            e.record()
        torch.cuda.synchronize()
        ms = sorted(s.elapsed_time(e) for s, e in ev)
        print(json.dumps({"emu_world": W, "T": T, "cfg": cfg, "ms_median": round(ms[len(ms) // 2], 4),
                          "ms_min": round(ms[0], 4), "kernel": plane.last()}), flush=True)
        plane.check()
        if dbg & 128:
            tr = plane.ws_trace(0).cpu()
            print_trace(tr)
    plane.close()


if __name__ == "__main__":
    main()
