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


def body_moves(comm, n):
    """allgatherMap / reduceScatterMap / gatherMap / broadcastMap / scatterMap with String keys and double values
    (PCS:666, 1299, 582, 842, 949): seconds per collective on this rank."""
    from ytk_mp4j_b200 import Operands, Operators
    r, p = comm.getRank(), comm.getSlaveNum()
    dop = Operands.DOUBLE_OPERAND()
    out = {}
    mine = {str(i): float(i) for i in range(r * n, (r + 1) * n)}
    lists = [{str(i): 1.0 for i in range(q * n, (q + 1) * n)} for q in range(p)]

    def timed(name, fn):
        comm.barrier()
        t0 = time.perf_counter()
        res = fn()
        out[name] = time.perf_counter() - t0
        return res
    res = timed("allgatherMap", lambda: comm.allgatherMap(mine, dop))
    assert len(res) == p and len(res[(r + 1) % p]) == n
    res = timed("reduceScatterMap", lambda: comm.reduceScatterMap(lists, dop, Operators.Double.SUM))
    assert len(res) == n and res[str(r * n)] == p
    res = timed("gatherMap", lambda: comm.gatherMap(mine, dop, 0))
    assert r != 0 or len(res) == n * p
    res = timed("broadcastMap", lambda: comm.broadcastMap(mine if r == 0 else {}, dop, 0))
    assert len(res) == n
    res = timed("scatterMap", lambda: comm.scatterMap(lists if r == 0 else None, dop, 0))
    assert len(res) == n
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--procs", type=int, default=4)
    ap.add_argument("--keys", default="1000,100000,1000000")
    ap.add_argument("--moves", action="store_true", help="time the five other map collectives instead")
    a = ap.parse_args()
    from ytk_mp4j_b200.parallel.local_cluster import run_cluster
    if a.moves:
        for n in [int(x) for x in a.keys.split(",")]:
            for native in ("1", "0"):
                res, code, _ = run_cluster(a.procs, body_moves, (n,), timeout=900, env={"MP4J_NATIVE_MAPS": native})
                row = {k: round(max(r[k] for r in res) * 1e3, 1) for k in res[0]}
                print(json.dumps({"bench": "host_map_collectives_string_keys_ms", "procs": a.procs, "keys_per_rank": n,
                                  "codec": "packed + native merge" if native == "1" else "pickle + python merge", **row}))
        return
    for n in [int(x) for x in a.keys.split(",")]:
        res, code, _ = run_cluster(a.procs, body, (n,), timeout=600)
        print(json.dumps({"bench": "host_allreduce_map_string_keys", "procs": a.procs, "keys_per_rank": n,
                          "ms": round(max(res) * 1e3, 2), "Mkeys_per_s": round(n / max(res) / 1e6, 3)}))


if __name__ == "__main__":
    main()
