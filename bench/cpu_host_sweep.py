#!/usr/bin/env python
"""Host-plane sweep on the CPU box: allreduce / reduce-scatter / allgather of doubles, shared-memory
windows (MP4J_SHM=1, ranks share the machine) against the TCP ring algorithms (MP4J_SHM=0, the
reference's algorithms and the multi-host path). Times only the collective (median of reps).

  python bench/cpu_host_sweep.py [--procs 2,4,8] [--elems 1,1000,100000,1000000,10000000]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402


def body(comm, sizes, reps):
    from ytk_mp4j_b200 import CommUtils, Operands, Operators
    op, S = Operands.DOUBLE_OPERAND(), Operators.Double.SUM
    p = comm.getSlaveNum()
    out = {"transport": comm.host_transport}
    for n in sizes:
        arr = np.ones(max(n, 1), np.float64)
        froms, tos = CommUtils.createProcessArrayFroms(n, p), CommUtils.createProcessArrayTos(n, p)
        counts = [t - f for f, t in zip(froms, tos)]
        for name, fn in (("allreduce", lambda: comm.allreduceArray(arr, op, S, 0, n)),
                         ("reduce_scatter", lambda: comm.reduceScatterArray(arr, op, S, 0, counts)),
                         ("allgather", lambda: comm.allgatherArray(arr, op, froms, tos))):
            r = max(3, min(reps, int(2e8 / max(n * 8, 1)) + 3))
            fn()
            ts = []
            for _ in range(r):
                comm.barrier()
                t0 = time.perf_counter()
                fn()
                ts.append(time.perf_counter() - t0)
            out["%s/%d" % (name, n)] = sorted(ts)[len(ts) // 2] * 1e6
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--procs", default="2,4,8")
    ap.add_argument("--elems", default="1,1000,100000,1000000,10000000")
    ap.add_argument("--reps", type=int, default=50)
    a = ap.parse_args()
    from ytk_mp4j_b200.parallel.local_cluster import run_cluster
    sizes = [int(x) for x in a.elems.split(",")]
    rows = []
    for p in [int(x) for x in a.procs.split(",")]:
        per = {}
        for shm in ("1", "0"):
            os.environ["MP4J_SHM"] = shm
            res, code, _ = run_cluster(p, body, (sizes, a.reps), timeout=900)
            assert code == 0
            tr = res[0]["transport"]
            for k in res[0]:
                if k != "transport":
                    per.setdefault(k, {})[tr] = max(r[k] for r in res)  # slowest rank
        for k, v in per.items():
            coll, n = k.split("/")
            rows.append({"procs": p, "collective": coll, "elems": int(n), "shm_us": round(v.get("shm", 0), 1),
                         "tcp_us": round(v.get("tcp", 0), 1),
                         "speedup": round(v["tcp"] / v["shm"], 2) if v.get("shm") and v.get("tcp") else None})
    for r in rows:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
