#!/usr/bin/env python
"""Generates docs/API.md from the package's docstrings and the C headers: python docs/gen_api.py"""
import inspect
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from ytk_mp4j_b200 import CommMaster, CommUtils, Operands, Operators, ProcessCommSlave, ThreadCommSlave  # noqa: E402
from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane, GpuPlane  # noqa: E402
from ytk_mp4j_b200.utils.scatter_allocate import ScatterAllocate  # noqa: E402

# one-liners for methods whose semantics are exactly the reference's and carry no docstring of their own
FALLBACK = {
    "gatherArray": "rank i's [from_i, to_i) lands at the same offsets of the root's array (in place)",
    "allgatherArray": "every rank ends with all p segments (in place)",
    "broadcastArray": "root's [from, to) to everybody",
    "scatterArray": "rank i receives the root's [from_i, to_i) at the same offsets",
    "reduceScatterArray": "rank r ends with the fully reduced [from + sum(counts[<r]), + counts[r])",
    "reduceArray": "root ends with the reduction of [from, to) over all participants",
    "allreduceArray": "everybody ends with the reduction of [from, to) over all participants",
    "allreduceArrayRpc": "small arrays: whole arrays meet (master mailbox on the host path, one-shot kernel on the GPU) and are folded in rank order",
    "broadcast": "scalar form of broadcastArray", "reduce": "scalar form of reduceArray",
    "allreduce": "scalar form of allreduceArray", "allreduceRpc": "scalar form of allreduceArrayRpc",
    "gatherMap": "root receives the union of all maps",
    "broadcastMap": "root's map to everybody", "scatterMap": "rank i receives map i of the root's list",
    "reduceScatterMap": "list[i] = contribution to rank i; returns the merged map this rank owns",
    "reduceMap": "root receives the key-wise reduction of all maps",
    "allreduceMap": "everybody receives the key-wise reduction of all maps (keys placed by String.hashCode() % p)",
    "barrier": "all participants meet (master barrier; thread barrier around it in ThreadCommSlave)",
    "threadBarrier": "the threads of this process meet", "close": "reports the exit code to the master and leaves the job",
    "info": "log line forwarded to the master (default: only rank 0 / thread 0)", "debug": "as info, debug level",
    "error": "error line forwarded to the master", "exception": "forwards the stack trace, then close(1)",
    "writeFile": "creates a text file ON THE MASTER", "getRank": "rank of this process", "getSlaveNum": "number of processes",
    "getThreadNum": "threads per process", "setThreadId": "binds the calling thread to thread id t", "getThreadId": "id of the calling thread (default 0)",
    "isClosed": "true after close()", "start": "starts the master service", "shutdown": "stops the master immediately",
    "logs": "log lines collected from the slaves", "port": "port the master listens on", "main": "command-line entry point / self-check",
}
FALLBACK.update({
    "gpu": "the GPU data plane of this communicator (created on first use)", "init_gpu": "collective set-up of the GPU plane: symmetric heap, peer mapping, signal pads, multicast",
    "rank": "rank of this process", "size": "number of processes", "allgatherMap": "list of the p maps indexed by origin rank",
    "BF16_OPERAND": "bfloat16 tensors (addition for torch users)", "HALF_OPERAND": "float16 tensors (addition for torch users)",
    "BYTE_OPERAND": "byte / int8 payloads", "SHORT_OPERAND": "short / int16 payloads", "INT_OPERAND": "int / int32 payloads",
    "LONG_OPERAND": "long / int64 payloads", "FLOAT_OPERAND": "float32 payloads; wire='bf16'|'fp8' opts into the lossy GPU wire codec",
    "DOUBLE_OPERAND": "float64 payloads", "STRING_OPERAND": "str payloads (host path)", "OBJECT_OPERAND": "arbitrary objects (pickle or a user codec; host path)",
    "createProcessArrayFroms": "balanced partition: start of every rank's slice (last rank takes the remainder)",
    "createProcessArrayTos": "balanced partition: end of every rank's slice", "createThreadArrayFroms": "balanced [slaveNum][threadNum] partition, starts",
    "createThreadArrayTos": "balanced [slaveNum][threadNum] partition, ends", "getFromsFromCount": "starts from a start offset and per-rank counts",
    "getTosFromCount": "ends from a start offset and per-rank counts", "getProcessFroms": "collapses 2-D thread ranges to per-process starts",
    "getProcessTos": "collapses 2-D thread ranges to per-process ends", "isFromCountsLegal": "raises unless from >= 0 and all counts >= 0",
    "isFromToLegal": "raises unless 0 <= from <= to", "isfromsTosLegal": "raises unless the ranges are ordered and non-overlapping (1-D or [slaveNum][threadNum])",
    "splitRange": "the balanced split the compositions use (PCS:1741-1756)", "recvNum": "number of messages a rank receives under a scatter plan",
    "allocate": "binomial scatter plan rooted at `root`", "allocateHalving": "the reference's seed-and-halve scatter plan", "split": "halving schedule of a rank range",
    "allgather": "dense allgather kernel on CUDA tensors", "gather": "dense gather kernel", "scatter": "dense scatter kernel",
    "reduce_scatter": "dense reduce-scatter kernel", "thread_allreduce": "fused thread-reduce + inter-GPU allreduce + fan-out kernel over T tensors",
    "describe": "mode, world, heap and multicast facts of the plane", "has_multicast": "True when NVLS (multimem) is available",
    "last": "name, algorithm, bytes and grid of the last collective", "world": "number of ranks of the plane",
    "allreduce_wire": "fp32 SUM allreduce with bf16 / block-scaled fp8 payload on the wire (lossy, opt-in)", "empty": "symmetric-heap tensors (zero-copy collectives)",
    "gbdt_hist_reduce_scatter": "K13 on emulated ranks: one (bins, grad, hess) set per rank", "lr_xtr_reduce_scatter": "K12 on emulated ranks: one (Xt, Rt) pair per rank",
    "check": "raises if a kernel hit the peer-wait watchdog", "close": "releases the plane",
})
for kind, what in (("SetUnion", "set union"), ("SetIntersection", "set intersection"), ("ListConcat", "list concatenation")):
    for pre in ("reduce", "allreduce"):
        FALLBACK[pre + kind] = "%s of one %s per participant (%s)" % (what, what.split()[0], "to the root" if pre == "reduce" else "to everybody")
        FALLBACK[pre + "Map" + kind] = "key-wise %s of maps of %ss (%s)" % (what, what.split()[0], "to the root" if pre == "reduce" else "to everybody")


def first(doc):
    return " ".join(l.strip() for l in doc.strip().split("\n\n")[0].split("\n")) if doc else ""


def dump(out, cls, title):
    out.append("## %s\n" % title)
    if cls.__doc__:
        out.append(first(cls.__doc__) + "\n")
    out += ["| method | summary |", "|---|---|"]
    names = [n for n, _ in inspect.getmembers(cls) if not n.startswith("_")]
    camel = {n.replace("_", "").lower() for n in names if "_" not in n}
    for name in names:
        if "_" in name and name.replace("_", "").lower() in camel:
            continue  # snake_case alias of a camelCase method
        raw = inspect.getattr_static(cls, name)
        target = raw.fget if isinstance(raw, property) else getattr(cls, name)
        if not callable(target):
            continue
        try:
            sig = str(inspect.signature(target))
        except (TypeError, ValueError):
            sig = "(...)"
        doc = first(getattr(target, "__doc__", None))
        if not doc:
            base = name[:-7] if name.endswith("Process") else name
            doc = FALLBACK.get(base, "")
            if name.endswith("Process") and doc:
                doc = "pass-through to the process communicator (one thread per process calls it): " + doc
        out.append("| `%s%s` | %s |" % (name, sig.replace("|", "\\|"), doc.replace("|", "\\|")))
    out.append("")


def main():
    out = ["# API reference (generated by `python docs/gen_api.py` from the docstrings and `include/*.h`)\n",
           "Method names and argument order follow the reference (`com.fenbi.mp4j.*`); file:line parity notes are in the "
           "docstrings. CUDA tensors take the GPU kernels, numpy / CPU tensors the native host plane (shared-memory windows "
           "on one machine, TCP otherwise), lists / dicts / Python operators the generic path. The Java facade (`java/`) "
           "exposes the same names over JNI.\n"]
    for cls, title in ((CommMaster, "CommMaster"), (ProcessCommSlave, "ProcessCommSlave"), (ThreadCommSlave, "ThreadCommSlave"),
                       (Operands, "Operands"), (CommUtils, "CommUtils"), (ScatterAllocate, "ScatterAllocate"),
                       (GpuPlane, "GpuPlane (`comm.gpu`)"), (EmuPlane, "EmuPlane (emulated ranks on one GPU, for tests)")):
        dump(out, cls, title)
    out.append("## Operators\n")
    for g in ("Double", "Float", "Long", "Int", "Short", "Byte"):
        grp = getattr(Operators, g)
        out.append("* `Operators.%s`: " % g + ", ".join("`%s`" % n for n in dir(grp) if not n.startswith("_")))
    out.append("* `Operators.custom(fn, name, vectorized)`: Python callable, host path. "
               "`Operators.device_expr(expr, name, host_fn)`: C++ expression over `a`, `b`, compiled with NVRTC into the "
               "fused GPU kernels.\n")
    out.append("## C ABI\n")
    for h in ("include/mp4j.h", "include/mp4j_gpu.h"):
        src = open(os.path.join(ROOT, h)).read()
        out += ["### %s\n" % h, "```c"]
        for m in re.finditer(r"^(?:int|void|const char\*|uint64_t|double|float|int32_t|int64_t|void\*)\s+\*?mp4j_\w+\([^;]*\);", src, re.M):
            out.append(re.sub(r"\s+", " ", m.group(0)))
        out.append("```\n")
    with open(os.path.join(ROOT, "docs", "API.md"), "w") as f:
        f.write("\n".join(out))
    print("docs/API.md: %d lines" % len(out))


if __name__ == "__main__":
    main()
