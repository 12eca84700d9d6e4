#!/usr/bin/env python
"""BASELINE.json config 1: "CommMaster + 4 ProcessCommSlave on localhost CPU: allreduceArray double
SUM, 1M elements". The reference's JVM cannot run on this image (no JDK), so this times the same
algorithms (ring reduce-scatter + ring allgather over loopback TCP, master barrier) in our C++ host
plane against the reference's published 1 GbE numbers. Runs on the CPU box.

  python bench/cpu_host_bench.py [--procs 4] [--elems 1000000] [--reps 20]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

PUBLISHED_MS = {  # allreduce double SUM, README.md:306-352: {slaves: {elements: ms}}
    2: {100000: 12, 1000000: 110, 10000000: 955}, 4: {100000: 23, 1000000: 130, 10000000: 1230},
    8: {100000: 28, 1000000: 183, 10000000: 1641},
}


def body(comm, n, reps):
    from ytk_mp4j_b200 import Operands, Operators
    op, S = Operands.DOUBLE_OPERAND(), Operators.Double.SUM
    arr = np.ones(n, np.float64)
    out = {}
    for name, fn in (("allreduce", lambda: comm.allreduceArray(arr, op, S, 0, n)),
                     ("allreduce_rpc", lambda: comm.allreduceArrayRpc(arr, op, S))):
        arr[:] = 1.0
        fn()
        comm.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            arr[:] = 1.0
            fn()
            comm.barrier()  # the reference's timing includes the trailing master barrier (PCS:1367)
        out[name] = (time.perf_counter() - t0) / reps * 1e3
    assert np.all(arr == comm.getSlaveNum())
    out["transport"] = comm.host_transport
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--procs", type=int, default=4)
    ap.add_argument("--elems", type=int, default=1000000)
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    from ytk_mp4j_b200.parallel.local_cluster import run_cluster
    res, code, _ = run_cluster(a.procs, body, (a.elems, a.reps), timeout=600)
    ms = max(r["allreduce"] for r in res)
    rpc = max(r["allreduce_rpc"] for r in res)
    pub = PUBLISHED_MS.get(a.procs, {}).get(a.elems)
    p = a.procs
    print(json.dumps({"bench": "host_allreduce_double_sum", "procs": p, "elems": a.elems, "ms": round(ms, 3),
                      "rpc_ms": round(rpc, 3), "busbw_GBps": round(a.elems * 8 / ms / 1e6 * 2 * (p - 1) / p, 3),
                      "published_reference_ms_1GbE": pub, "speedup_vs_published": round(pub / ms, 1) if pub else None,
                      "transport": {"shm": "shared-memory windows, one gather-reduce-scatter pass (C++ host plane)",
                                    "tcp": "loopback TCP, ring RS + ring AG (C++ host plane)"}[res[0]["transport"]],
                      "master_exit_code": code}))


if __name__ == "__main__":
    main()
