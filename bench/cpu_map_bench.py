#!/usr/bin/env python
"""Host allreduceMap with String keys (the reference's map path, PCS:2053-2081) on local ranks.
  python bench/cpu_map_bench.py [--procs 4] [--keys 1000,100000,1000000]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def body(comm, n):
    from ytk_mp4j_b200 import Operands, Operators
    r = comm.getRank()
    m = {str(i): 1.0 for i in range(n)}
    m[str(-(r + 1))] = 1.0  # shared keys + one private key per rank (ProcessAllReduceCheck.java:86-116)
    comm.barrier()
    t0 = time.perf_counter()
    out = comm.allreduceMap(dict(m), Operands.DOUBLE_OPERAND(), Operators.Double.SUM)
    dt = time.perf_counter() - t0
    assert len(out) == n + comm.getSlaveNum() and out["5"] == comm.getSlaveNum()
    return dt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--procs", type=int, default=4)
    ap.add_argument("--keys", default="1000,100000,1000000")
    a = ap.parse_args()
    from ytk_mp4j_b200.parallel.local_cluster import run_cluster
    for n in [int(x) for x in a.keys.split(",")]:
        res, code, _ = run_cluster(a.procs, body, (n,), timeout=600)
        print(json.dumps({"bench": "host_allreduce_map_string_keys", "procs": a.procs, "keys_per_rank": n,
                          "ms": round(max(res) * 1e3, 2), "Mkeys_per_s": round(n / max(res) / 1e6, 3)}))


if __name__ == "__main__":
    main()
