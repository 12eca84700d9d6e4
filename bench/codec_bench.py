#!/usr/bin/env python
"""K8: fp32 SUM allreduce with the payload crossing NVLink as bf16 / block-scaled fp8 (lossy, opt-in) against the
lossless two-shot P2P kernel and AUTO (NVLS where available) on REAL links. Run under torchrun.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/codec_bench.py"""
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from ytk_mp4j_b200 import CommMaster, Operands, Operators, ProcessCommSlave  # noqa: E402
from ytk_mp4j_b200 import _native as N  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    lr = int(os.environ.get("LOCAL_RANK", 0))
    port = int(os.environ.get("MASTER_PORT", 29500)) + 37
    torch.cuda.set_device(lr)
    master = CommMaster(world, port, quiet=True, work_dir=tempfile.gettempdir()).start() if rank == 0 else None
    comm = ProcessCommSlave("codec", "127.0.0.1", port, rank_hint=rank, host="127.0.0.1")
    plane = comm.init_gpu(device=lr, heap_bytes=3 << 30)
    fop, SUM = Operands.FLOAT_OPERAND(), Operators.Float.SUM
    dop, DMAX = Operands.DOUBLE_OPERAND(), Operators.Double.MAX
    for nbytes in (16 << 20, 256 << 20, 1 << 30):
        n = nbytes // 4
        x = comm.empty(n, fop)
        x.fill_(0.001)
        cases = {
            "lossless two-shot": lambda: plane.allreduce(x, fop, SUM, 0, n, algo=N.ALGO_TWOSHOT),
            "lossless auto": lambda: plane.allreduce(x, fop, SUM, 0, n),
            "wire bf16": lambda: plane._allreduce_wire([x], 0, n, N.WIRE_BF16, None),
            "wire fp8": lambda: plane._allreduce_wire([x], 0, n, N.WIRE_FP8, None),
        }
        row = {"bench": "allreduce_wire", "n_gpus": world, "bytes": nbytes}
        for name, fn in cases.items():
            x.fill_(0.001)
            for _ in range(2):
                fn()
            torch.cuda.synchronize()
            comm.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                fn()
            e1.record()
            torch.cuda.synchronize()
            us = comm.allreduce(e0.elapsed_time(e1) * 1e3 / 5, dop, DMAX)
            row[name + " us"] = round(us, 1)
            row[name + " busbw GB/s"] = round(nbytes / us / 1e3 * 2 * (world - 1) / world, 1)
        if rank == 0:
            print(json.dumps(row), flush=True)
        plane.free(x)
    plane.check()
    comm.close(0)
    if master:
        master.stop(timeout_s=10)
        master.close()


if __name__ == "__main__":
    main()
