#!/usr/bin/env python
"""Hierarchical host allreduce on one box: MP4J_NODE_ID splits the local ranks into "machines" (the network between
them is loopback, i.e. as fast as memory - the case that favours the flat ring most). Alternates the hierarchical
form (per-machine shared memory + TCP between machines) with the flat TCP ring (MP4J_HIER=0).

  python bench/cpu_hier_bench.py [--ids 0,0,1,1] [--elems 10000000] [--rounds 3]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def body(comm, n, reps):
    import time
    import numpy as np
    from ytk_mp4j_b200 import Operands, Operators
    op, S = Operands.DOUBLE_OPERAND(), Operators.Double.SUM
    arr = np.ones(n)
    comm.allreduceArray(arr, op, S, 0, n)
    ts = []
    for _ in range(reps):
        comm.barrier()
        t0 = time.perf_counter()
        comm.allreduceArray(arr, op, S, 0, n)
        ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3, comm.host_transport


def main():
    from ytk_mp4j_b200.parallel.local_cluster import run_cluster
    ap = argparse.ArgumentParser()
    ap.add_argument("--ids", default="0,0,1,1", help="machine id of every rank")
    ap.add_argument("--elems", type=int, default=10_000_000)
    ap.add_argument("--rounds", type=int, default=3)
    a = ap.parse_args()
    p = len(a.ids.split(","))
    for rnd in range(a.rounds):
        for hier in ("1", "0"):
            res, code, _ = run_cluster(p, body, (a.elems, 8), timeout=600, env={"MP4J_NODE_IDS": a.ids, "MP4J_HIER": hier})
            print(json.dumps({"bench": "host_allreduce_double_sum", "machines": a.ids, "elems": a.elems,
                              "transport": res[0][1], "ms_best_of_8": round(max(r[0] for r in res), 1)}), flush=True)


if __name__ == "__main__":
    main()
