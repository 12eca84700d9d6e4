"""The self-check matrix (parity: the 128 classes of J/check/check{double,float,long,int,short,byte,
string,object}/{Process,Thread}{Gather,Scatter,Allgather,ReduceScatter,Broadcast,Reduce,AllReduce,
RpcAllReduce}Check.java, bases J/check/{CommCheck,ProcessCheck,ThreadCheck}.java).

The reference copy-pastes one class per (type, level, collective); here one parametrised function
per collective covers all eight operand types, both levels (process / thread) and - new - the
device ("cpu" numpy arrays or "cuda" tensors). Every check follows the reference pattern:
deterministic fill -> time the call -> info("... takes: N ms") -> element-wise verify -> on mismatch
info(..., onlyRank0=False) and close(1) (e.g. checkdouble/ProcessAllReduceCheck.java:61-72).
"""
import time

import numpy as np

from .. import CommUtils, Operands, Operators
from ..operators import custom

PRIMS = {
    "double": (Operands.DOUBLE_OPERAND, Operators.Double, np.float64),
    "float": (Operands.FLOAT_OPERAND, Operators.Float, np.float32),
    "long": (Operands.LONG_OPERAND, Operators.Long, np.int64),
    "int": (Operands.INT_OPERAND, Operators.Int, np.int32),
    "short": (Operands.SHORT_OPERAND, Operators.Short, np.int16),
    "byte": (Operands.BYTE_OPERAND, Operators.Byte, np.int8),
}
TYPES = list(PRIMS) + ["string", "object"]
COLLECTIVES = ["gather", "scatter", "allgather", "reducescatter", "broadcast", "reduce", "allreduce"]


class CheckFailed(Exception):
    pass


class _Ctx:
    """What a check needs to know about where it runs (process or thread level)."""

    def __init__(self, comm, tid=None):
        self.comm, self.tid = comm, tid
        self.thread = tid is not None
        self.p = comm.getSlaveNum()
        self.r = comm.getRank()
        self.T = comm.getThreadNum() if self.thread else 1
        self.P = self.p * self.T               # participants
        self.g = self.r * self.T + (tid or 0)  # global participant id

    def info(self, msg):
        self.comm.info(msg)

    def fail(self, msg):
        if self.thread:
            self.comm.info(msg, False)
        else:
            self.comm.info(msg, False)
        raise CheckFailed(msg)


def _make(dtype, n, value, device):
    if dtype in PRIMS:
        arr = np.full(n, value, PRIMS[dtype][2])
        if device == "cuda":
            import torch
            return torch.from_numpy(arr).cuda()
        return arr
    if dtype == "string":
        return [str(value)] * n
    return [{"val": value} for _ in range(n)]  # ObjectNode{val} (checkobject/ObjectNode.java:31-44)


def _set(arr, f, t, value, dtype):
    if dtype in PRIMS:
        arr[f:t] = value
    elif dtype == "string":
        arr[f:t] = [str(value)] * (t - f)
    else:
        arr[f:t] = [{"val": value} for _ in range(t - f)]


def _vals(arr, dtype):
    if dtype in PRIMS:
        a = arr.cpu().numpy() if hasattr(arr, "cpu") else arr
        return a.astype(np.int64)
    if dtype == "string":
        return np.array([int(x) for x in arr], dtype=np.int64)
    return np.array([x["val"] for x in arr], dtype=np.int64)


def _operand_op(dtype, compress):
    if dtype in PRIMS:
        mk, ops, _ = PRIMS[dtype]
        return mk(compress), ops.SUM
    if dtype == "string":  # checkstring/ProcessAllReduceCheck.java:65-70
        return Operands.STRING_OPERAND(compress), custom(lambda a, b: str(int(a) + int(b)), "STR_SUM")
    return Operands.OBJECT_OPERAND(compress=compress), custom(lambda a, b: {"val": a["val"] + b["val"]}, "NODE_SUM")


def _wrap(dtype, v):
    """value of participant-count sums after the operand's narrowing (short/byte wrap)."""
    if dtype in PRIMS:
        return int(np.array(v).astype(PRIMS[dtype][2], casting="unsafe"))
    return v


def _ranges(ctx, n):
    if ctx.thread:
        f2, t2 = CommUtils.createThreadArrayFroms(n, ctx.p, ctx.T), CommUtils.createThreadArrayTos(n, ctx.p, ctx.T)
        return f2, t2, f2[ctx.r][ctx.tid], t2[ctx.r][ctx.tid]
    f, t = CommUtils.createProcessArrayFroms(n, ctx.p), CommUtils.createProcessArrayTos(n, ctx.p)
    return f, t, f[ctx.r], t[ctx.r]


def _owner_pattern(ctx, n, dtype):
    f, t, _, _ = _ranges(ctx, n)
    exp = np.full(n, -1, np.int64)
    for q in range(ctx.p):
        if ctx.thread:
            for th in range(ctx.T):
                exp[f[q][th]:t[q][th]] = _wrap(dtype, q * ctx.T + th)
        else:
            exp[f[q]:t[q]] = _wrap(dtype, q)
    return exp


def _timed(ctx, name, dtype, fn):
    start = time.time()
    out = fn()
    if hasattr(out, "is_cuda") and out.is_cuda:
        import torch
        torch.cuda.synchronize()
    ctx.info("%s %s %s takes:%d" % ("thread" if ctx.thread else "process", dtype, name,
                                    int((time.time() - start) * 1000)))
    return out


def check_collective(ctx, coll, dtype, n, compress=False, device="cpu", root=0, root_thread=0):
    """One (collective, type) cell of the matrix on arrays; raises CheckFailed on a wrong result."""
    c, T = ctx.comm, ctx.thread
    operand, op = _operand_op(dtype, compress)
    f, t, myf, myt = _ranges(ctx, n)
    is_root = ctx.r == root and (not T or ctx.tid == root_thread)
    rt = (root, root_thread) if T else (root,)
    owner = _owner_pattern(ctx, n, dtype)
    me = _wrap(dtype, ctx.g)
    if coll == "gather":
        arr = _make(dtype, n, -1, device)
        _set(arr, myf, myt, me, dtype)
        _timed(ctx, coll, dtype, lambda: c.gatherArray(arr, operand, f, t, *rt))
        if is_root and not np.array_equal(_vals(arr, dtype), owner):
            ctx.fail("gather %s array error" % dtype)
    elif coll == "allgather":
        arr = _make(dtype, n, -1, device)
        _set(arr, myf, myt, me, dtype)
        _timed(ctx, coll, dtype, lambda: c.allgatherArray(arr, operand, f, t))
        if not np.array_equal(_vals(arr, dtype), owner):
            ctx.fail("allgather %s array error" % dtype)
    elif coll == "scatter":
        arr = _make(dtype, n, -1, device)
        if is_root:
            for q in range(ctx.p):
                for th in range(ctx.T):
                    lo, hi = (f[q][th], t[q][th]) if T else (f[q], t[q])
                    _set(arr, lo, hi, _wrap(dtype, q * ctx.T + th), dtype)
        _timed(ctx, coll, dtype, lambda: c.scatterArray(arr, operand, f, t, *rt))
        if not np.all(_vals(arr, dtype)[myf:myt] == me):
            ctx.fail("scatter %s array error" % dtype)
    elif coll == "reducescatter":
        arr = _make(dtype, n, 0, device)
        for q in range(ctx.p):
            for th in range(ctx.T):
                lo, hi = (f[q][th], t[q][th]) if T else (f[q], t[q])
                _set(arr, lo, hi, _wrap(dtype, q * ctx.T + th), dtype)
        counts = [[t[q][th] - f[q][th] for th in range(ctx.T)] for q in range(ctx.p)] if T else \
            [t[q] - f[q] for q in range(ctx.p)]
        _timed(ctx, coll, dtype, lambda: c.reduceScatterArray(arr, operand, op, 0, counts))
        if not np.all(_vals(arr, dtype)[myf:myt] == _wrap(dtype, me * ctx.P)):
            ctx.fail("reduce scatter %s array error" % dtype)
    elif coll == "broadcast":
        arr = _make(dtype, n, 1 if is_root else -1, device)
        _timed(ctx, coll, dtype, lambda: c.broadcastArray(arr, operand, 0, n, *rt))
        if not np.all(_vals(arr, dtype) == 1):
            ctx.fail("broadcast %s array error" % dtype)
    elif coll == "reduce":
        arr = _make(dtype, n, 1, device)
        _timed(ctx, coll, dtype, lambda: c.reduceArray(arr, operand, op, 0, n, *rt))
        if is_root and not np.all(_vals(arr, dtype) == _wrap(dtype, ctx.P)):
            ctx.fail("reduce %s array error" % dtype)
    elif coll == "allreduce":
        arr = _make(dtype, n, 1, device)
        _timed(ctx, coll, dtype, lambda: c.allreduceArray(arr, operand, op, 0, n))
        if not np.all(_vals(arr, dtype) == _wrap(dtype, ctx.P)):
            ctx.fail("allreduce %s array error" % dtype)
    else:
        raise ValueError(coll)


def check_map(ctx, dtype, obj, compress=False, root=0, root_thread=0):
    """Map container cells (e.g. checkdouble/ProcessAllReduceCheck.java:86-116)."""
    c, T = ctx.comm, ctx.thread
    operand, op = _operand_op(dtype, compress)
    rt = (root, root_thread) if T else (root,)
    is_root = ctx.r == root and (not T or ctx.tid == root_thread)

    def val(v):
        return str(v) if dtype == "string" else ({"val": v} if dtype == "object" else PRIMS[dtype][2](v).item())

    def num(v):
        return int(v) if dtype == "string" else (v["val"] if dtype == "object" else int(v))
    m = {str(i): val(1) for i in range(obj)}
    m[str(-(ctx.g + 1))] = val(1)
    out = _timed(ctx, "allreduce map", dtype, lambda: c.allreduceMap(m, operand, op))
    if len(out) != obj + ctx.P or any(num(out[str(i)]) != _wrap(dtype, ctx.P) for i in range(obj)):
        ctx.fail("allreduce %s map error" % dtype)
    m = {str(i): val(i % 100) for i in range(ctx.g * obj, (ctx.g + 1) * obj)}
    out = _timed(ctx, "gather map", dtype, lambda: c.gatherMap(m, operand, *rt))
    if is_root and len(out) != ctx.P * obj:
        ctx.fail("gather %s map error" % dtype)
    out = _timed(ctx, "reduce map", dtype,
                 lambda: c.reduceMap({str(i): val(1) for i in range(obj)}, operand, op, *rt))
    if is_root and any(num(v) != _wrap(dtype, ctx.P) for v in out.values()):
        ctx.fail("reduce %s map error" % dtype)
    out = _timed(ctx, "broadcast map", dtype,
                 lambda: c.broadcastMap({str(i): val(i % 100) for i in range(obj)} if is_root else {}, operand, *rt))
    if len(out) != obj:
        ctx.fail("broadcast %s map error" % dtype)
    # scatterMap: the root holds one map per participant, participant g receives map g
    # (checkdouble/ProcessScatterCheck.java:100-131; thread level: [slaveNum][threadNum], TCS:1097)
    def part(g):
        return {str(j): val(j % 100) for j in range(g * obj, (g + 1) * obj)}
    if T:
        src = [[part(q * ctx.T + th) for th in range(ctx.T)] for q in range(ctx.p)] if is_root else None
    else:
        src = [part(q) for q in range(ctx.p)] if is_root else None
    out = _timed(ctx, "scatter map", dtype, lambda: c.scatterMap(src, operand, *rt))
    if len(out) != obj or any(num(out.get(str(j), val(-1))) != _wrap(dtype, j % 100)
                              for j in range(ctx.g * obj, (ctx.g + 1) * obj)):
        ctx.fail("scatter %s map error" % dtype)
    # allgatherMap: everybody ends with the list of all processes' maps
    # (checkdouble/ProcessAllgatherCheck.java:91-119; the T maps of a process are merged first, TCS:706)
    out = _timed(ctx, "allgather map", dtype, lambda: c.allgatherMap(part(ctx.g), operand))
    if len(out) != ctx.p:
        ctx.fail("allgather %s map error: %d maps" % (dtype, len(out)))
    for q in range(ctx.p):
        for g in range(q * ctx.T, (q + 1) * ctx.T):
            if any(num(out[q].get(str(j), val(-1))) != _wrap(dtype, j % 100) for j in range(g * obj, (g + 1) * obj)):
                ctx.fail("allgather %s map error" % dtype)
    # reduceScatterMap: everybody contributes one map per participant, participant g gets map g reduced
    # (checkdouble/ProcessReduceScatterCheck.java:93-122)
    def ones(g):
        return {str(j): val(1) for j in range(g * obj, (g + 1) * obj)}
    src = [[ones(q * ctx.T + th) for th in range(ctx.T)] for q in range(ctx.p)] if T else [ones(q) for q in range(ctx.p)]
    out = _timed(ctx, "reducescatter map", dtype, lambda: c.reduceScatterMap(src, operand, op))
    if len(out) != obj or any(num(out.get(str(j), val(0))) != _wrap(dtype, ctx.P)
                              for j in range(ctx.g * obj, (ctx.g + 1) * obj)):
        ctx.fail("reducescatter %s map error" % dtype)


def check_rpc(ctx, dtype, n, run_time, compress=False):
    """ring allreduce vs RPC allreduce, timed over run_time repetitions
    (checkdouble/ProcessRpcAllReduceCheck.java:56-110)."""
    c = ctx.comm
    operand, op = _operand_op(dtype, compress)
    for name, fn in (("allreduce", lambda a: c.allreduceArray(a, operand, op, 0, n)),
                     ("rpc allreduce", lambda a: c.allreduceArrayRpc(a, operand, op))):
        start = time.time()
        for _ in range(run_time):
            arr = _make(dtype, n, 1, "cpu")
            fn(arr)
            if not np.all(_vals(arr, dtype) == _wrap(dtype, ctx.P)):
                ctx.fail("%s %s error" % (name, dtype))
        ctx.info("%s %s x%d takes:%d" % (dtype, name, run_time, int((time.time() - start) * 1000)))


def check_sets(ctx, obj, root=0, root_thread=0):
    """set-union / set-intersection / list-concat helpers (checkobject/ProcessAllReduceCheck.java:139-257)."""
    c, T = ctx.comm, ctx.thread
    s = set(range(obj)) | {-(ctx.g + 1)}
    if len(c.allreduceSetUnion(set(s))) != obj + ctx.P:
        ctx.fail("allreduce set union error")
    inter = c.allreduceSetIntersection(set(s))
    if inter != (set(range(obj)) if ctx.P > 1 else s):
        ctx.fail("allreduce set intersection error")
    if sorted(c.allreduceListConcat([ctx.g])) != list(range(ctx.P)):
        ctx.fail("allreduce list concat error")
    rt = (root, root_thread) if T else (root,)
    u = c.reduceSetUnion(set(s), *rt)
    is_root = ctx.r == root and (not T or ctx.tid == root_thread)
    if is_root and len(u) != obj + ctx.P:
        ctx.fail("reduce set union error")
    # the other rooted helpers (checkobject/ProcessReduceCheck.java:147, 185, 234)
    inter = c.reduceSetIntersection(set(s), *rt)
    if is_root and inter != (set(range(obj)) if ctx.P > 1 else s):
        ctx.fail("reduce set intersection error")
    lst = c.reduceListConcat([ctx.g], *rt)
    if is_root and sorted(lst) != list(range(ctx.P)):
        ctx.fail("reduce list concat error")
    # the Map* forms: one set / list per key (checkobject/ProcessAllReduceCheck.java:139-257)
    keys = ["k%d" % i for i in range(3)]
    ms = {k: set(s) for k in keys}
    out = c.allreduceMapSetUnion({k: set(v) for k, v in ms.items()})
    if sorted(out) != keys or any(len(out[k]) != obj + ctx.P for k in keys):
        ctx.fail("allreduce map set union error")
    out = c.allreduceMapSetIntersection({k: set(v) for k, v in ms.items()})
    if any(out[k] != (set(range(obj)) if ctx.P > 1 else s) for k in keys):
        ctx.fail("allreduce map set intersection error")
    out = c.allreduceMapListConcat({k: [ctx.g] for k in keys})
    if any(sorted(out[k]) != list(range(ctx.P)) for k in keys):
        ctx.fail("allreduce map list concat error")
    out = c.reduceMapSetUnion({k: set(v) for k, v in ms.items()}, *rt)
    if is_root and any(len(out[k]) != obj + ctx.P for k in keys):
        ctx.fail("reduce map set union error")
    out = c.reduceMapSetIntersection({k: set(v) for k, v in ms.items()}, *rt)
    if is_root and any(out[k] != (set(range(obj)) if ctx.P > 1 else s) for k in keys):
        ctx.fail("reduce map set intersection error")
    out = c.reduceMapListConcat({k: [ctx.g] for k in keys}, *rt)
    if is_root and any(sorted(out[k]) != list(range(ctx.P)) for k in keys):
        ctx.fail("reduce map list concat error")


def check_process_passthrough(ctx, dtype, n, obj, compress=False):
    """The xxxProcess block of the thread checks (checkbyte/ThreadAllReduceCheck.java:160-240): ONE thread of every
    process calls the process-level collective through the ThreadCommSlave while the others wait at the thread
    barrier; the participants are then the slaveNum processes, not slaveNum x threadNum."""
    c = ctx.comm
    if not ctx.thread:
        return
    operand, op = _operand_op(dtype, compress)
    c.threadBarrier()
    if ctx.tid == 0:
        arr = _make(dtype, n, 1, "cpu")
        c.allreduceArrayProcess(arr, operand, op, 0, n)
        if not np.all(_vals(arr, dtype) == _wrap(dtype, ctx.p)):
            ctx.fail("allreduceArrayProcess %s error" % dtype)

        def val(v):
            return str(v) if dtype == "string" else ({"val": v} if dtype == "object" else PRIMS[dtype][2](v).item())

        def num(v):
            return int(v) if dtype == "string" else (v["val"] if dtype == "object" else int(v))
        m = {str(i): val(1) for i in range(obj)}
        m[str(-(ctx.r + 1))] = val(1)
        out = c.allreduceMapProcess(m, operand, op)
        if len(out) != obj + ctx.p or any(num(out[str(i)]) != _wrap(dtype, ctx.p) for i in range(obj)):
            ctx.fail("allreduceMapProcess %s error" % dtype)
        one = c.allreduceProcess(val(1), operand, op)
        if num(one) != _wrap(dtype, ctx.p):
            ctx.fail("allreduceProcess %s error" % dtype)
    c.threadBarrier()


def run_all(comm, tid, arr_size, obj_size, run_time, compress, test_rpc, device="cpu", types=None):
    """The body of CommCheckTool (J/check/CommCheckTool.java:57-105): every check, run_time times."""
    ctx = _Ctx(comm, tid)
    types = types or TYPES
    for _ in range(run_time):
        for dtype in types:
            if device == "cuda" and dtype not in PRIMS:
                continue
            n = arr_size if dtype in PRIMS else obj_size
            for coll in COLLECTIVES:
                check_collective(ctx, coll, dtype, n, compress, device)
            if device == "cpu":
                check_map(ctx, dtype, obj_size, compress)
                check_process_passthrough(ctx, dtype, n, obj_size, compress)
        if device == "cpu":
            check_sets(ctx, obj_size)
    if test_rpc and device == "cpu":
        for dtype in types:
            check_rpc(ctx, dtype, min(arr_size, 100000) if dtype in PRIMS else obj_size, run_time, compress)
    return True
