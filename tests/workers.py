"""Worker bodies for the multi-process tests (spawned by parallel.local_cluster.run_cluster).
Value patterns are the reference check harness's (J/check/check*/Process*Check.java,
Thread*Check.java), extended with what it never exercises: non-zero roots, from > 0, odd sizes,
every operator."""
import os

import numpy as np

from ytk_mp4j_b200 import CommUtils, Mp4jException, Operands, Operators
from ytk_mp4j_b200.operators import custom

DT = {
    "double": (Operands.DOUBLE_OPERAND, Operators.Double, np.float64),
    "float": (Operands.FLOAT_OPERAND, Operators.Float, np.float32),
    "long": (Operands.LONG_OPERAND, Operators.Long, np.int64),
    "int": (Operands.INT_OPERAND, Operators.Int, np.int32),
    "short": (Operands.SHORT_OPERAND, Operators.Short, np.int16),
    "byte": (Operands.BYTE_OPERAND, Operators.Byte, np.int8),
}


def _expected_owner(n, p):
    avg = n // p
    idx = np.arange(n)
    return np.minimum(idx // avg, p - 1) if avg > 0 else np.full(n, p - 1)


def process_arrays(comm, dtype, n, root, compress=False, transport=None):
    """All 7 array collectives + scalars for one primitive dtype through the C++ host plane."""
    if transport is not None and comm.getSlaveNum() > 1:
        assert comm.host_transport == transport, comm.host_transport
    mk, ops, npdt = DT[dtype]
    operand = mk(compress)
    p, r = comm.getSlaveNum(), comm.getRank()
    froms, tos = CommUtils.createProcessArrayFroms(n, p), CommUtils.createProcessArrayTos(n, p)
    owner = _expected_owner(n, p).astype(npdt)
    # gather: own slice = rank; root sees owner pattern (ProcessGatherCheck.java:54-84)
    arr = np.full(n, -1, npdt)
    arr[froms[r]:tos[r]] = r
    comm.gatherArray(arr, operand, froms, tos, root)
    if r == root:
        assert np.array_equal(arr, owner), "gather"
    # allgather (ProcessAllgatherCheck.java:55-83)
    arr = np.full(n, -1, npdt)
    arr[froms[r]:tos[r]] = r
    comm.allgatherArray(arr, operand, froms, tos)
    assert np.array_equal(arr, owner), "allgather"
    # scatter: root fills slice q with q, others -1 (ProcessScatterCheck.java:55-90)
    arr = owner.copy() if r == root else np.full(n, -1, npdt)
    comm.scatterArray(arr, operand, froms, tos, root)
    assert np.all(arr[froms[r]:tos[r]] == r), "scatter"
    # reduce-scatter: arr[i] = i/avg on every rank -> own slice = owner*p (ProcessReduceScatterCheck.java:56-85)
    arr = owner.copy()
    counts = [t - f for f, t in zip(froms, tos)]
    comm.reduceScatterArray(arr, operand, ops.SUM, 0, counts)
    assert np.array_equal(arr[froms[r]:tos[r]], (owner[froms[r]:tos[r]].astype(np.int64) * p).astype(npdt)), "rs"
    # broadcast with from > 0 (ProcessBroadcastCheck.java:59-77 uses from = 0 only)
    lo = min(3, n)
    arr = np.full(n, 1 if r == root else -1, npdt)
    comm.broadcastArray(arr, operand, lo, n, root)
    assert np.all(arr[lo:] == 1), "broadcast"
    if r != root:
        assert np.all(arr[:lo] == -1), "broadcast must not touch [0, from)"
    assert comm.broadcast(npdt(r).item(), operand, root) == root, "scalar broadcast"
    # reduce (ProcessReduceCheck.java:58-75)
    arr = np.ones(n, npdt)
    comm.reduceArray(arr, operand, ops.SUM, 0, n, root)
    if r == root:
        assert np.all(arr == p), "reduce"
    v = comm.reduce(npdt(1).item(), operand, ops.SUM, root)
    if r == root:
        assert v == p
    # allreduce: all ones -> p (ProcessAllReduceCheck.java:61-72), with from > 0 as well
    arr = np.ones(n, npdt)
    comm.allreduceArray(arr, operand, ops.SUM, 0, n)
    assert np.all(arr == p), "allreduce"
    arr = np.ones(n, npdt)
    comm.allreduceArray(arr, operand, ops.SUM, lo, n)
    assert np.all(arr[lo:] == p) and np.all(arr[:lo] == 1), "allreduce from>0"
    assert comm.allreduce(npdt(1).item(), operand, ops.SUM) == p
    # MAX / MIN / PROD never run end-to-end in the reference
    arr = np.full(n, r + 1, npdt)
    comm.allreduceArray(arr, operand, ops.MAX, 0, n)
    assert np.all(arr == p), "max"
    arr = np.full(n, r + 1, npdt)
    comm.allreduceArray(arr, operand, ops.MIN, 0, n)
    assert np.all(arr == 1), "min"
    arr = np.full(n, 2 if r == 0 else 1, npdt)
    comm.allreduceArray(arr, operand, ops.PROD, 0, n)
    assert np.all(arr == 2), "prod"
    if dtype in ("long", "int", "short", "byte"):
        arr = np.full(n, 1 << r, npdt)
        comm.allreduceArray(arr, operand, ops.BITS_OR, 0, n)
        assert np.all(arr == (1 << p) - 1), "or"
        arr = np.full(n, 7, npdt)
        comm.allreduceArray(arr, operand, ops.BITS_AND, 0, n)
        assert np.all(arr == 7), "and"
        arr = np.full(n, 1, npdt)
        comm.allreduceArray(arr, operand, ops.BITS_XOR, 0, n)
        assert np.all(arr == (p & 1)), "xor"
    # RPC allreduce (ProcessRpcAllReduceCheck.java:56-110)
    arr = np.ones(n, npdt)
    comm.allreduceArrayRpc(arr, operand, ops.SUM)
    assert np.all(arr == p), "rpc"
    assert comm.allreduceRpc(npdt(2).item(), operand, ops.SUM) == 2 * p
    comm.barrier()
    return True


def process_loc_ops(comm):
    p, r = comm.getSlaveNum(), comm.getRank()
    D, L = Operators.Double, Operators.Long
    arr = np.array([D.compositeDouble(float(r), r), D.compositeDouble(-float(r), r), D.compositeDouble(1.0, r)])
    comm.allreduceArray(arr, Operands.DOUBLE_OPERAND(), D.FLOAT_MAX_LOC, 0, 3)
    assert D.getIntLoc(arr[0]) == p - 1 and D.getFloatVal(arr[0]) == p - 1
    assert D.getIntLoc(arr[1]) == 0
    assert D.getIntLoc(arr[2]) == 0  # tie -> lowest loc
    arr = np.array([L.compositeLong(r, r), L.compositeLong(5, r)], dtype=np.int64)
    comm.allreduceArray(arr, Operands.LONG_OPERAND(), L.INT_MIN_LOC, 0, 2)
    assert L.getIntLoc(arr[0]) == 0 and L.getIntLoc(arr[1]) == 0
    return True


def process_generic(comm, obj_size, root, compress):
    """String / Object arrays, maps for every operand kind, custom operators, set/list helpers
    (J/check/checkstring, J/check/checkobject, J/check/checkdouble map parts)."""
    p, r = comm.getSlaveNum(), comm.getRank()
    n = obj_size
    froms, tos = CommUtils.createProcessArrayFroms(n, p), CommUtils.createProcessArrayTos(n, p)
    owner = [int(x) for x in _expected_owner(n, p)]
    sop = Operands.STRING_OPERAND(compress)
    str_sum = custom(lambda a, b: str(int(a) + int(b)), "STR_SUM")  # checkstring/ProcessAllReduceCheck.java:65-70
    # String arrays
    arr = ["-1"] * n
    arr[froms[r]:tos[r]] = [str(r)] * (tos[r] - froms[r])
    comm.gatherArray(arr, sop, froms, tos, root)
    if r == root:
        assert arr == [str(o) for o in owner]
    arr = ["-1"] * n
    arr[froms[r]:tos[r]] = [str(r)] * (tos[r] - froms[r])
    comm.allgatherArray(arr, sop, froms, tos)
    assert arr == [str(o) for o in owner]
    arr = [str(o) for o in owner] if r == root else ["-1"] * n
    comm.scatterArray(arr, sop, froms, tos, root)
    assert arr[froms[r]:tos[r]] == [str(r)] * (tos[r] - froms[r])
    arr = ["1"] * n
    comm.allreduceArray(arr, sop, str_sum, 0, n)
    assert arr == [str(p)] * n
    arr = ["1"] * n
    comm.reduceArray(arr, sop, str_sum, 0, n, root)
    if r == root:
        assert arr == [str(p)] * n
    arr = ["1" if r == root else "-1"] * n
    comm.broadcastArray(arr, sop, 0, n, root)
    assert arr == ["1"] * n
    arr = [str(o) for o in owner]
    comm.reduceScatterArray(arr, sop, str_sum, 0, [t - f for f, t in zip(froms, tos)])
    assert arr[froms[r]:tos[r]] == [str(r * p)] * (tos[r] - froms[r])
    assert comm.allreduce("2", sop, str_sum) == str(2 * p)
    assert comm.broadcast(str(r), sop, root) == str(root)
    arr = ["1"] * 5
    comm.allreduceArrayRpc(arr, sop, str_sum)
    assert arr == [str(p)] * 5
    # Object arrays: dicts standing in for ObjectNode{val} (checkobject/ObjectNode.java:31-44)
    oop = Operands.OBJECT_OPERAND(compress=compress)
    node_sum = custom(lambda a, b: {"val": a["val"] + b["val"]}, "NODE_SUM")
    arr = [{"val": 1} for _ in range(n)]
    comm.allreduceArray(arr, oop, node_sum, 0, n)
    assert all(x["val"] == p for x in arr)
    # numpy array + custom python operator (scalar and vectorised)
    a = np.ones(n, np.float64)
    comm.allreduceArray(a, Operands.DOUBLE_OPERAND(compress), custom(lambda x, y: x + y), 0, n)
    assert np.all(a == p)
    a = np.ones(n, np.float32)
    comm.allreduceArray(a, Operands.FLOAT_OPERAND(compress), custom(lambda x, y: x + y, vectorized=True), 0, n)
    assert np.all(a == p)
    # maps (ProcessAllReduceCheck.java:86-116 etc.)
    dop = Operands.DOUBLE_OPERAND(compress)
    m = {str(i): 1.0 for i in range(n)}
    m[str(-(r + 1))] = 1.0
    out = comm.allreduceMap(m, dop, Operators.Double.SUM)
    assert len(out) == n + p
    assert all(out[str(i)] == p for i in range(n)) and all(out[str(-(q + 1))] == 1 for q in range(p))
    m = {str(i): float(i) for i in range(r * n, (r + 1) * n)}
    out = comm.gatherMap(m, dop, root)
    if r == root:
        assert len(out) == p * n and all(out[str(i)] == i for i in range(p * n))
    lst = comm.allgatherMap({str(i): float(i) for i in range(r * n, (r + 1) * n)}, dop)
    assert len(lst) == p and all(sorted(int(k) for k in lst[q]) == list(range(q * n, (q + 1) * n)) for q in range(p))
    maps = [{str(i): float(i) for i in range(q * n, (q + 1) * n)} for q in range(p)] if r == root else None
    mine = comm.scatterMap(maps, dop, root)
    assert sorted(int(k) for k in mine) == list(range(r * n, (r + 1) * n))
    maps = [{str(i): 1.0 for i in range(q * n, (q + 1) * n)} for q in range(p)]
    mine = comm.reduceScatterMap(maps, dop, Operators.Double.SUM)
    assert len(mine) == n and all(v == p for v in mine.values())
    out = comm.broadcastMap({str(i): float(i) for i in range(n)} if r == root else {}, dop, root)
    assert len(out) == n and all(out[str(i)] == i for i in range(n))
    out = comm.reduceMap({str(i): 1.0 for i in range(n)}, dop, Operators.Double.SUM, root)
    if r == root:
        assert len(out) == n and all(v == p for v in out.values())
    out = comm.allreduceMap({"k%d" % i: "1" for i in range(n)}, sop, str_sum)
    assert all(v == str(p) for v in out.values()) and len(out) == n
    out = comm.allreduceMap({str(i): 1 for i in range(n)}, Operands.INT_OPERAND(compress), Operators.Int.MAX)
    assert all(v == 1 for v in out.values())
    # set / list helpers (checkobject/ProcessAllReduceCheck.java:139-257, ProcessReduceCheck.java:147-234)
    s = set(range(n)) | {-(r + 1)}
    u = comm.allreduceSetUnion(set(s))
    assert len(u) == n + p
    i = comm.allreduceSetIntersection(set(s))
    assert i == (set(range(n)) if p > 1 else s)
    c = comm.allreduceListConcat([r])
    assert sorted(c) == list(range(p))
    mu = comm.allreduceMapSetUnion({"a": {r}, "b": {0}})
    assert mu["a"] == set(range(p)) and mu["b"] == {0}
    mi = comm.allreduceMapSetIntersection({"a": {r, 100}})
    assert mi["a"] == ({100} if p > 1 else {r, 100})
    mc = comm.allreduceMapListConcat({"a": [r]})
    assert sorted(mc["a"]) == list(range(p))
    u = comm.reduceSetUnion(set(s), root)
    if r == root:
        assert len(u) == n + p
    i = comm.reduceSetIntersection(set(s), root)
    if r == root:
        assert i == (set(range(n)) if p > 1 else s)
    c = comm.reduceListConcat([r], root)
    if r == root:
        assert sorted(c) == list(range(p))
    comm.reduceMapSetUnion({"a": {r}}, root)
    comm.reduceMapSetIntersection({"a": {r}}, root)
    comm.reduceMapListConcat({"a": [r]}, root)
    comm.barrier()
    return True


def thread_arrays(comm, tid, dtype, n, root, root_thread):
    """Thread-level collectives on host arrays (J/check/checkdouble/Thread*Check.java)."""
    mk, ops, npdt = DT[dtype]
    operand = mk()
    p, r, T = comm.getSlaveNum(), comm.getRank(), comm.getThreadNum()
    g = r * T + tid  # global participant id
    f2, t2 = CommUtils.createThreadArrayFroms(n, p, T), CommUtils.createThreadArrayTos(n, p, T)
    expect = np.full(n, -1, npdt)
    for q in range(p):
        for t in range(T):
            expect[f2[q][t]:t2[q][t]] = q * T + t
    # allreduce: all ones -> p*T (ThreadAllReduceCheck.java:76-87)
    arr = np.ones(n, npdt)
    comm.allreduceArray(arr, operand, ops.SUM, 0, n)
    assert np.all(arr == p * T), "thread allreduce"
    # gather with 2-D ranges (ThreadGatherCheck.java:73-95)
    arr = np.full(n, -1, npdt)
    arr[f2[r][tid]:t2[r][tid]] = g
    comm.gatherArray(arr, operand, f2, t2, root, root_thread)
    if r == root and tid == root_thread:
        assert np.array_equal(arr, expect), "thread gather"
    # allgather
    arr = np.full(n, -1, npdt)
    arr[f2[r][tid]:t2[r][tid]] = g
    comm.allgatherArray(arr, operand, f2, t2)
    assert np.array_equal(arr, expect), "thread allgather"
    # scatter (ThreadScatterCheck.java:71-94)
    arr = expect.copy() if (r == root and tid == root_thread) else np.full(n, -1, npdt)
    comm.scatterArray(arr, operand, f2, t2, root, root_thread)
    assert np.all(arr[f2[r][tid]:t2[r][tid]] == g), "thread scatter"
    # reduce-scatter with 2-D counts (ThreadReduceScatterCheck.java:72-105)
    counts = [[t2[q][t] - f2[q][t] for t in range(T)] for q in range(p)]
    arr = expect.copy()
    comm.reduceScatterArray(arr, operand, ops.SUM, 0, counts)
    want = (expect[f2[r][tid]:t2[r][tid]].astype(np.int64) * p * T).astype(npdt)
    assert np.array_equal(arr[f2[r][tid]:t2[r][tid]], want), "thread reduce-scatter"
    # broadcast
    arr = np.full(n, 1 if (r == root and tid == root_thread) else -1, npdt)
    comm.broadcastArray(arr, operand, 0, n, root, root_thread)
    assert np.all(arr == 1), "thread broadcast"
    # reduce
    arr = np.ones(n, npdt)
    comm.reduceArray(arr, operand, ops.SUM, 0, n, root, root_thread)
    if r == root and tid == root_thread:
        assert np.all(arr == p * T), "thread reduce"
    # scalars
    assert comm.allreduce(npdt(1).item(), operand, ops.SUM) == p * T
    assert comm.allreduceRpc(npdt(1).item(), operand, ops.SUM) == p * T
    assert comm.broadcast(npdt(g % 100).item(), operand, root, root_thread) == (root * T + root_thread) % 100
    v = comm.reduce(npdt(1).item(), operand, ops.SUM, root, root_thread)
    assert v == (p * T if (r == root and tid == root_thread) else 1)
    arr = np.ones(n, npdt)
    comm.allreduceArrayRpc(arr, operand, ops.SUM)
    assert np.all(arr == p * T)
    # xxxProcess pass-through: one thread per process (checkbyte/Thread*Check extra block)
    comm.threadBarrier()
    if tid == 0:
        arr = np.ones(n, npdt)
        comm.allreduceArrayProcess(arr, operand, ops.SUM, 0, n)
        assert np.all(arr == p)
    comm.barrier()
    return True


def thread_maps(comm, tid, n, root, root_thread):
    p, r, T = comm.getSlaveNum(), comm.getRank(), comm.getThreadNum()
    g = r * T + tid
    dop = Operands.DOUBLE_OPERAND()
    S = Operators.Double.SUM
    m = {str(i): 1.0 for i in range(n)}
    m[str(-(g + 1))] = 1.0
    out = comm.allreduceMap(m, dop, S)
    assert len(out) == n + p * T and all(out[str(i)] == p * T for i in range(n))
    out = comm.gatherMap({str(i): float(i) for i in range(g * n, (g + 1) * n)}, dop, root, root_thread)
    if r == root and tid == root_thread:
        assert len(out) == p * T * n
    else:
        assert out is None
    lst = comm.allgatherMap({str(i): float(i) for i in range(g * n, (g + 1) * n)}, dop)
    assert len(lst) == p and all(len(lst[q]) == T * n for q in range(p))
    out = comm.broadcastMap({"a": 1.0} if (r == root and tid == root_thread) else {}, dop, root, root_thread)
    assert out == {"a": 1.0}
    mll = None
    if r == root and tid == root_thread:
        mll = [[{str(q * T + t): 1.0} for t in range(T)] for q in range(p)]
    mine = comm.scatterMap(mll, dop, root, root_thread)
    assert mine == {str(g): 1.0}
    mll = [[{str(q * T + t): 1.0} for t in range(T)] for q in range(p)]
    mine = comm.reduceScatterMap(mll, dop, S)
    assert mine == {str(g): float(p * T)}
    out = comm.reduceMap({str(i): 1.0 for i in range(n)}, dop, S, root, root_thread)
    if r == root and tid == root_thread:
        assert all(v == p * T for v in out.values()) and len(out) == n
    else:
        assert out is None
    u = comm.allreduceSetUnion({g})
    assert u == set(range(p * T))
    c = comm.allreduceListConcat([g])
    assert sorted(c) == list(range(p * T))
    i = comm.allreduceSetIntersection({g, 1000})
    assert i == ({1000} if p * T > 1 else {g, 1000})
    u = comm.reduceSetUnion({g}, root, root_thread)
    if r == root and tid == root_thread:
        assert u == set(range(p * T))
    comm.barrier()
    return True


def control_plane(comm):
    p, r = comm.getSlaveNum(), comm.getRank()
    comm.info("hello from everyone", onlyRank0=False)
    comm.info("hello from rank0 only")
    comm.debug("dbg", onlyRank0=False)
    comm.error("an error line")
    comm.barrier()
    if r == 0:
        comm.writeFile("content-%d" % p, "mp4j_writefile_test.txt")
    comm.barrier()
    return comm.getRank()


def failing_slave(comm):
    if comm.getRank() == 1:
        try:
            raise ValueError("boom")
        except ValueError as e:
            comm.exception(e, sleep_s=0.1)
        return "closed-1"
    return "ok"


def dying_peer(comm, n=1000):
    """Rank 1 vanishes without closing; rank 0's next collective must fail instead of hanging."""
    import os
    import time
    arr = np.ones(n)
    comm.allreduceArray(arr, Operands.DOUBLE_OPERAND(), Operators.Double.SUM, 0, n)  # everybody is set up
    if comm.getRank() == 1:
        os._exit(7)
    t0 = time.time()
    try:
        comm.allreduceArray(arr, Operands.DOUBLE_OPERAND(), Operators.Double.SUM, 0, n)
    except Mp4jException as e:
        return "failed-fast: %s (%.1f s)" % (str(e)[:80], time.time() - t0)
    return "no error"


def native_maps(comm):
    """String-keyed maps with primitive values through the C++ map collective (csrc/host/map_host.cc):
    non-ASCII keys (UTF-16 hashCode placement), empty contributions, non-zero root, several value types."""
    p, r = comm.getSlaveNum(), comm.getRank()
    dop, lop = Operands.DOUBLE_OPERAND(), Operands.LONG_OPERAND()
    keys = ["k%d" % i for i in range(500)] + ["été-%d" % i for i in range(50)] + ["漢字%d" % i for i in range(50)] \
        + ["\U0001F600%d" % i for i in range(20)] + [""]
    m = {k: float(i % 7 + 1) for i, k in enumerate(keys)}
    m["private-%d" % r] = 100.0 + r
    out = comm.allreduceMap(dict(m), dop, Operators.Double.SUM)
    assert len(out) == len(keys) + p, (len(out), len(keys) + p)
    for i, k in enumerate(keys):
        assert out[k] == p * float(i % 7 + 1), (k, out[k])
    for q in range(p):
        assert out["private-%d" % q] == 100.0 + q
    # only even ranks contribute; MAX over longs; result on the last rank only
    contrib = {("x%d" % i): (r + 1) * 1000 + i for i in range(300)} if r % 2 == 0 else {}
    root = p - 1
    out = comm.reduceMap(contrib, lop, Operators.Long.MAX, root)
    top = max(q for q in range(p) if q % 2 == 0)
    if r == root:
        assert len(out) == 300 and all(out["x%d" % i] == (top + 1) * 1000 + i for i in range(300)), len(out)
    else:
        assert out == {}
    out = comm.allreduceMap({}, dop, Operators.Double.SUM)
    assert out == {}
    try:
        comm.allreduceMap({1: 2.0}, dop, Operators.Double.SUM)
        assert p == 1 or os.environ.get("MP4J_NATIVE_MAPS") == "0", "int keys must be rejected on the native path"
    except Mp4jException:
        pass
    return True


def map_bridge_worker(comm, mask):
    """String-key bridge of the GPU map path, host part only (no GPU needed): identical dictionaries on every
    rank, round trip, and the collision policy (a narrowed hash makes two strings share a key)."""
    from ytk_mp4j_b200.comm import map_bridge
    map_bridge._TEST_MASK = mask
    r, p = comm.getRank(), comm.getSlaveNum()
    br = map_bridge.MapKeyBridge(comm)
    shared = ["f%d" % i for i in range(200)]
    mine = ["rank%d_%d" % (r, i) for i in range(50)] + ["ключ-%d" % r, "鍵%d" % r]
    ks = br.encode(shared + mine)
    everything = shared + [s for q in range(p) for s in ["rank%d_%d" % (q, i) for i in range(50)] + ["ключ-%d" % q, "鍵%d" % q]]
    out = {"broken": br.broken, "n": len(br._rev)}
    if not br.broken:
        allk = map_bridge.keys64(everything)
        out["roundtrip"] = br.decode(allk.tolist()) == everything          # other ranks' strings are known here
        out["mine"] = br.decode(ks.tolist()) == shared + mine
        ks2 = br.encode(shared + mine)                                      # nothing new: no string traffic
        out["stable"] = bool((ks2 == ks).all()) and len(br._rev) == out["n"]
        tagged = map_bridge.MapKeyBridge.tag(ks, 5)
        out["tag"] = bool(((tagged.astype("uint64") >> 60) == 5).all()) and br.decode(tagged.tolist(), True) == shared + mine
    return out


def plane_blocker(comm, off_rank):
    """What every rank concludes about the GPU plane when ONE rank sets MP4J_GPU_PLANE=0 (off_rank < 0: nobody)."""
    import os
    if comm.getRank() == off_rank:
        os.environ["MP4J_GPU_PLANE"] = "0"
    else:
        os.environ.pop("MP4J_GPU_PLANE", None)
    return comm._gpu_plane_blocker()


def ragged_arrays(comm, seed, n):
    """All seven array collectives on doubles with RAGGED per-rank ranges (empty and very unequal slices, gaps between
    them, from > 0) against numpy: what balanced splits never exercise - with a small shared-memory window the
    multi-pass part arithmetic of the node plane, with MP4J_SHM=0 the ring's uneven segments."""
    import numpy as np
    from ytk_mp4j_b200 import Operands, Operators
    p, r = comm.getSlaveNum(), comm.getRank()
    rnd = np.random.RandomState(seed)  # the same stream on every rank
    dop = Operands.DOUBLE_OPERAND()
    ok = True
    for it in range(6):
        cuts = np.sort(rnd.randint(0, n, size=2 * p))
        froms, tos = [int(x) for x in cuts[0::2]], [int(x) for x in cuts[1::2]]
        k = int(rnd.randint(p))
        tos[k] = froms[k]  # an empty slice
        data = [rnd.randint(-50, 50, size=n).astype(np.float64) for _ in range(p)]
        root = int(rnd.randint(p))
        # allgather / gather / scatter
        a = data[r].copy()
        comm.allgatherArray(a, dop, froms, tos)
        want = data[r].copy()
        for q in range(p):
            want[froms[q]:tos[q]] = data[q][froms[q]:tos[q]]
        ok &= np.array_equal(a, want)
        a = data[r].copy()
        comm.gatherArray(a, dop, froms, tos, root)
        if r == root:
            ok &= np.array_equal(a, want)
        a = data[r].copy()
        comm.scatterArray(a, dop, froms, tos, root)
        # only the own slice is specified: a rank that relays in the scatter tree holds the slices it forwards
        # at their offsets afterwards (as in the reference)
        ok &= np.array_equal(a[froms[r]:tos[r]], data[root][froms[r]:tos[r]])
        # reduce-scatter with ragged counts, allreduce / reduce / broadcast on a sub-range
        counts = [int(x) for x in rnd.randint(0, n // p, size=p)]
        counts[int(rnd.randint(p))] = 0
        f0 = int(rnd.randint(0, n - sum(counts) + 1))
        total = sum(data)
        a = data[r].copy()
        comm.reduceScatterArray(a, dop, Operators.Double.SUM, f0, counts)
        lo = f0 + sum(counts[:r])
        ok &= np.array_equal(a[lo:lo + counts[r]], total[lo:lo + counts[r]])
        f, t = sorted(int(x) for x in rnd.randint(0, n + 1, size=2))
        a = data[r].copy()
        comm.allreduceArray(a, dop, Operators.Double.MAX, f, t)
        w3 = data[r].copy()
        w3[f:t] = np.maximum.reduce(data)[f:t]
        ok &= np.array_equal(a, w3)
        a = data[r].copy()
        comm.reduceArray(a, dop, Operators.Double.SUM, f, t, root)
        if r == root:
            ok &= np.array_equal(a[f:t], total[f:t]) and np.array_equal(a[:f], data[r][:f]) and np.array_equal(a[t:], data[r][t:])
        a = data[r].copy()
        comm.broadcastArray(a, dop, f, t, root)
        w4 = data[r].copy()
        w4[f:t] = data[root][f:t]
        ok &= np.array_equal(a, w4)
        if not ok:
            return "mismatch in iteration %d" % it
    return True


def ragged_thread_arrays(comm, tid, seed, n):
    """ThreadCommSlave's seven array collectives with ragged [slaveNum][threadNum] ranges / counts (empty slices,
    gaps, non-zero roots) on doubles against numpy - the native mp4j::ThreadComm paths."""
    import numpy as np
    from ytk_mp4j_b200 import Operands, Operators
    p, r, T = comm.getSlaveNum(), comm.getRank(), comm.getThreadNum()
    rnd = np.random.RandomState(seed)  # the same stream on every thread of every rank
    dop = Operands.DOUBLE_OPERAND()
    bad = []  # never leave early: the other threads would wait for this one in the next collective
    for it in range(5):
        cuts = np.sort(rnd.randint(0, n, size=2 * p * T))
        lo = [[int(cuts[2 * (q * T + t)]) for t in range(T)] for q in range(p)]
        hi = [[int(cuts[2 * (q * T + t) + 1]) for t in range(T)] for q in range(p)]
        eq, et = int(rnd.randint(p)), int(rnd.randint(T))
        hi[eq][et] = lo[eq][et]  # an empty slice
        data = [[rnd.randint(-50, 50, size=n).astype(np.float64) for _ in range(T)] for _ in range(p)]
        root, rt = int(rnd.randint(p)), int(rnd.randint(T))
        mine = data[r][tid]
        gathered = mine.copy()
        for q in range(p):
            for t in range(T):
                gathered[lo[q][t]:hi[q][t]] = data[q][t][lo[q][t]:hi[q][t]]
        # only the slices are specified: between the thread slices of a process the reference ships whatever its
        # delegate thread holds (process ranges are [lo[q][0], hi[q][T-1]), TCS:621-676), the GPU kernel ships nothing
        def slices_ok(x):
            return all(np.array_equal(x[lo[q][t]:hi[q][t]], gathered[lo[q][t]:hi[q][t]]) for q in range(p) for t in range(T))
        a = mine.copy()
        comm.allgatherArray(a, dop, lo, hi)
        if not slices_ok(a):
            bad.append("allgather %d" % it)
        a = mine.copy()
        comm.gatherArray(a, dop, lo, hi, root, rt)
        if r == root and tid == rt and not slices_ok(a):
            bad.append("gather %d" % it)
        a = mine.copy()
        comm.scatterArray(a, dop, lo, hi, root, rt)
        if not np.array_equal(a[lo[r][tid]:hi[r][tid]], data[root][rt][lo[r][tid]:hi[r][tid]]):
            bad.append("scatter %d" % it)
        counts = [[int(x) for x in rnd.randint(0, n // (p * T), size=T)] for _ in range(p)]
        counts[int(rnd.randint(p))][int(rnd.randint(T))] = 0
        tot_cnt = sum(sum(c) for c in counts)
        f0 = int(rnd.randint(0, n - tot_cnt + 1))
        total = sum(sum(row) for row in data)
        a = mine.copy()
        comm.reduceScatterArray(a, dop, Operators.Double.SUM, f0, counts)
        off = f0 + sum(sum(c) for c in counts[:r]) + sum(counts[r][:tid])
        if not np.array_equal(a[off:off + counts[r][tid]], total[off:off + counts[r][tid]]):
            bad.append("reduce-scatter %d" % it)
        f, t = sorted(int(x) for x in rnd.randint(0, n + 1, size=2))
        a = mine.copy()
        comm.allreduceArray(a, dop, Operators.Double.SUM, f, t)
        if not (np.array_equal(a[f:t], total[f:t]) and np.array_equal(a[:f], mine[:f]) and np.array_equal(a[t:], mine[t:])):
            bad.append("allreduce %d" % it)
        a = mine.copy()
        comm.reduceArray(a, dop, Operators.Double.MAX, f, t, root, rt)
        if r == root and tid == rt and not np.array_equal(a[f:t], np.maximum.reduce([x for row in data for x in row])[f:t]):
            bad.append("reduce %d" % it)
        a = mine.copy()
        comm.broadcastArray(a, dop, f, t, root, rt)
        if not np.array_equal(a[f:t], data[root][rt][f:t]):
            bad.append("broadcast %d" % it)
    return bad or True


def many_slaves(comm):
    """Barrier + allreduce in a job with more ranks than cores (the reference runs on clusters of tens of slaves)."""
    import numpy as np
    from ytk_mp4j_b200 import Operands, Operators
    for _ in range(5):
        comm.barrier()
    a = np.arange(5003, dtype=np.float64)
    comm.allreduceArray(a, Operands.DOUBLE_OPERAND(), Operators.Double.SUM, 0, 5003)
    m = comm.allreduceMap({"k%d" % comm.getRank(): 1, "all": 1}, Operands.INT_OPERAND(), Operators.Int.SUM)
    return bool(np.array_equal(a, np.arange(5003) * comm.getSlaveNum())) and m["all"] == comm.getSlaveNum() \
        and len(m) == comm.getSlaveNum() + 1


def hierarchical(comm, n):
    """Several "machines" (MP4J_NODE_IDS) with several ranks each: allreduce runs reduce-scatter in shared memory,
    TCP between the machines, allgather in shared memory; everything else stays on the TCP mesh."""
    import numpy as np
    from ytk_mp4j_b200 import Operands, Operators, CommUtils
    p, r = comm.getSlaveNum(), comm.getRank()
    assert comm.host_transport == "tcp+shm", comm.host_transport
    dop = Operands.DOUBLE_OPERAND()
    rnd = np.random.RandomState(99)
    data = [rnd.randint(-50, 50, size=n).astype(np.float64) for _ in range(p)]
    ok = True
    for f, t in ((0, n), (3, n - 5), (7, 9), (n // 2, n // 2)):
        a = data[r].copy()
        comm.allreduceArray(a, dop, Operators.Double.SUM, f, t)
        want = data[r].copy()
        want[f:t] = sum(data)[f:t]
        ok &= bool(np.array_equal(a, want))
        a = data[r].copy()
        comm.allreduceArray(a, dop, Operators.Double.MIN, f, t)
        want = data[r].copy()
        want[f:t] = np.minimum.reduce(data)[f:t]
        ok &= bool(np.array_equal(a, want))
    # the other collectives still work (flat TCP)
    froms, tos = CommUtils.createProcessArrayFroms(n, p), CommUtils.createProcessArrayTos(n, p)
    a = data[r].copy()
    comm.allgatherArray(a, dop, froms, tos)
    want = np.concatenate([data[q][froms[q]:tos[q]] for q in range(p)])
    ok &= bool(np.array_equal(a, want))
    a = data[r].copy()
    comm.reduceScatterArray(a, dop, Operators.Double.SUM, 0, [t - f for f, t in zip(froms, tos)])
    ok &= bool(np.array_equal(a[froms[r]:tos[r]], sum(data)[froms[r]:tos[r]]))
    comm.barrier()
    return ok
