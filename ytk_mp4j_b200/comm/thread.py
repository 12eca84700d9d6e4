"""ThreadCommSlave: one communicator shared by T threads of a process; every collective exists
as `xxx(...)` (all slaveNum x threadNum threads participate) and `xxxProcess(...)` (one thread
per process calls; pass-through to ProcessCommSlave). Parity: J/comm/ThreadCommSlave.java.

Design (not a translation): the reference pairs threads through an Exchanger and reduces in
lg T full-length passes (TCS:259-303), then thread 0 runs the process collective and the others
copy the result out behind two barriers (TCS:1941-1957). Here the T threads deposit their
arguments in per-thread slots, meet at a native spin-then-sleep barrier (csrc/host/thread_group.cc)
and ONE delegate thread does the work for all:
  * CUDA tensors and pinned host tensors: every thread calls into the C ABI with its own buffer; the rendezvous, the
    launch by one thread and the delivery of its error to all threads are native (mp4j::ThreadComm::delegate,
    csrc/host/thread_comm.h). The launch is a single fused kernel that reads the T thread buffers, reduces them in
    registers, runs the inter-GPU step over NVLink and fans the result out - never T x the data on the wire;
  * native host arrays: mp4j::ThreadComm, all T threads combine and fan out in parallel;
  * Python objects (maps, scalars, lists): `_delegate` below - the delegate reduces/merges in thread-id order, calls
    the process collective and hands results back.
Root-thread delegation follows SURVEY §2.7-14: only the root process honours rootThreadId, every
other process delegates through thread 0; non-root threads keep their own input (arrays) or get
None (maps). Map results are shared by reference among the threads of a process (TCS:2214-2217).
"""
import ctypes as C
import threading

import numpy as np

from .. import _native as N
from ..exception import Mp4jException, check
from ..utils.comm_utils import CommUtils
from .process import ProcessCommSlave, _is_cuda, _is_torch

# mp4j_collective codes of include/mp4j_types.h (native thread-level collectives)
_GATHER, _SCATTER, _ALL_GATHER, _BROADCAST, _REDUCE_SCATTER, _REDUCE = 0, 1, 2, 3, 4, 5


class ThreadCommSlave:
    def __init__(self, loginName, threadNum, masterHost, masterPort, rank_hint=None, host=None):
        if threadNum < 1 or threadNum > N.MAX_THREADS:
            raise Mp4jException("threadNum must be in [1, %d]" % N.MAX_THREADS)
        self.processCommSlave = ProcessCommSlave(loginName, masterHost, masterPort, rank_hint, host)
        self._T = int(threadNum)
        self._tg = C.c_void_p()
        check(N.lib.mp4j_tgroup_create(self._T, C.byref(self._tg)))
        self._tls = threading.local()
        self._slots = [None] * self._T
        self._result = [None, None]
        self._exc = [None, None]

    # ------------------------------------------------------------------ identity
    def getThreadNum(self):
        return self._T

    def isClosed(self):
        return self.processCommSlave.isClosed()

    def setThreadId(self, threadId):
        if threadId < 0 or threadId >= self._T:
            raise Mp4jException("threadId out of range")
        self._tls.tid = int(threadId)

    def getThreadId(self):
        return getattr(self._tls, "tid", 0)  # defaults to 0 (TCS:119-125)

    def getRank(self):
        return self.processCommSlave.getRank()

    def getSlaveNum(self):
        return self.processCommSlave.getSlaveNum()

    @property
    def gpu(self):
        return self.processCommSlave.gpu

    def init_gpu(self, *a, **kw):
        return self.processCommSlave.init_gpu(*a, **kw)

    def _cuda_plane(self):
        """The GPU plane for a thread-level collective on CUDA tensors, or None when the job cannot have one (ranks
        on several machines: ProcessCommSlave._gpu_plane_blocker). Whether it can, and setting it up, are collective
        over the ranks, so ONE thread does that for the process: on its first CUDA collective every thread passes
        through this rendezvous once; afterwards it is a field read."""
        pcs = self.processCommSlave
        if not getattr(self._tls, "plane_known", False):
            self.threadBarrier()
            if self.getThreadId() == 0 and pcs._gpu is None:
                try:
                    pcs._cuda_plane()
                    self._plane_exc = None
                except BaseException as e:  # noqa: BLE001 - delivered to every thread below
                    self._plane_exc = e
            self.threadBarrier()
            self._tls.plane_known = True
            e = getattr(self, "_plane_exc", None)
            if e is not None:
                raise e if isinstance(e, Mp4jException) else Mp4jException(repr(e))
        return pcs._gpu

    # ------------------------------------------------------------------ sync / logging
    def threadBarrier(self):
        check(N.lib.mp4j_tgroup_barrier(self._tg))

    def barrier(self):
        """thread barrier -> thread 0 master barrier -> thread barrier (TCS:151-157)."""
        self.threadBarrier()
        if self.getThreadId() == 0:
            self.processCommSlave.barrier()
        self.threadBarrier()

    def info(self, info, onlyRank0Thread0=True):
        if onlyRank0Thread0:
            if self.getThreadId() == 0:
                self.processCommSlave.info(info, True)
        else:
            self.processCommSlave.info("[threadId=%d] %s" % (self.getThreadId(), info), False)  # TCS:198

    def debug(self, debug, onlyRank0Thread0=True):
        if onlyRank0Thread0:
            if self.getThreadId() == 0:
                self.processCommSlave.debug(debug, True)
        else:
            self.processCommSlave.debug("[threadId=%d] %s" % (self.getThreadId(), debug), False)

    def error(self, error):
        self.processCommSlave.error("[threadId=%d] %s" % (self.getThreadId(), error))

    def exception(self, e, sleep_s=5.0):
        self.processCommSlave.exception(e, sleep_s)

    def close(self, code=0):
        self.processCommSlave.close(code)

    def __del__(self):
        try:
            if self._tg:
                N.lib.mp4j_tgroup_destroy(self._tg)
                self._tg = C.c_void_p()
        except Exception:
            pass

    # ------------------------------------------------------------------ delegate machinery
    def _delegate(self, my_args, fn, root_thread=0):
        """All T threads call with their own arguments; `root_thread` runs fn(args_by_thread) and
        its return value (a list indexed by thread id, or a single shared object) reaches
        everyone. Two barriers per collective; results are double-buffered by call parity."""
        tid = self.getThreadId()
        par = getattr(self._tls, "calls", 0) & 1
        self._tls.calls = getattr(self._tls, "calls", 0) + 1
        self._slots[tid] = my_args
        self.threadBarrier()
        if tid == root_thread:
            try:
                self._result[par] = fn(list(self._slots))
                self._exc[par] = None
            except BaseException as e:  # surfaced on every thread
                self._result[par] = None
                self._exc[par] = e
        self.threadBarrier()
        if self._exc[par] is not None:
            e = self._exc[par]
            raise e if isinstance(e, Mp4jException) else Mp4jException(repr(e))
        return self._result[par]

    def _native_thread(self, kind, arrData, operand, operator, from_=0, to=0, lo=None, hi=None, counts=None,
                       rootRank=0, rootThreadId=0):
        """Host arrays with a native dtype (and a predefined operator where one applies): the C++ thread
        collective in which all T threads combine and fan out in parallel (mp4j_thread_collective).
        Returns False if this call must take the generic delegate path instead."""
        if _is_cuda(arrData):
            return False
        a = self.processCommSlave._native_host(arrData, operand, operator)
        if a is None:
            return False

        def flat(m):
            return N.i64_array([x for row in m for x in row]) if m is not None else None
        check(N.lib.mp4j_thread_collective(self._tg, self.processCommSlave._h, self.getThreadId(), kind, a.ctypes.data,
                                           operand.code, operator.code if operator is not None else N.COPY, from_, to,
                                           flat(lo), flat(hi), flat(counts), rootRank, rootThreadId))
        return True

    def _gpu_segments(self, name, arrData, operand, operator, segs, root_thread=0):
        """CUDA tensors: the whole collective (thread step + inter-GPU step + thread step) as ONE launch of the
        segment kernel (csrc/gpu/thread_seg.cu). Returns False if this call cannot take it (too many segments,
        custom operator, 16-bit floats): the caller then composes it from separate steps."""
        if not _is_cuda(arrData):
            return False
        segs = [s for s in segs if s[1] > s[0]]
        plane = self._cuda_plane()
        if plane is None or not plane.thread_segments_ok([arrData], operand, operator, len(segs)):
            return False

        plane.thread_native_segments(self._tg, self.getThreadId(), root_thread, name, arrData, operand, operator, segs)
        return True

    def _real_root(self, rootRank, rootThreadId):
        if rootThreadId < 0 or rootThreadId >= self._T:
            raise Mp4jException("rootThreadId out of range")
        return rootThreadId if self.getRank() == rootRank else 0  # TCS:486

    # local (intra-process) combine helpers -----------------------------------------------
    def _host(self, arr, operand):
        return self.processCommSlave._generic_array(arr, operand)

    def _local_reduce(self, arrs, operand, operator, f, t, dst):
        """arrs[dst][f:t] = op over threads in thread-id order (thread path, DOP:441-457)."""
        if f >= t or len(arrs) == 1:
            return
        if _is_cuda(arrs[0]):
            plane = self.processCommSlave._gpu
            if plane is not None:
                plane.thread_reduce(arrs, operand, operator, f, t, dst)
            else:
                self._torch_reduce(arrs, operand, operator, f, t, dst)
            return
        order = [dst] + [i for i in range(len(arrs)) if i != dst]
        acc = self._host(arrs[dst], operand)
        native = operand.is_primitive and operator.is_native and isinstance(acc, np.ndarray)
        merge = None if native else ProcessCommSlave._merge_seq(operator, operand)
        for i in order[1:]:
            src = self._host(arrs[i], operand)
            if native:
                seg = np.ascontiguousarray(src[f:t])
                check(N.lib.mp4j_host_reduce_into(operand.code, operator.code,
                                                  acc[f:t].ctypes.data, seg.ctypes.data, t - f))
            else:
                acc[f:t] = merge(acc[f:t], src[f:t])

    def _torch_reduce(self, arrs, operand, operator, f, t, dst):
        """Thread reduce of CUDA tensors in a job without a GPU plane (ranks on several machines): torch ops for
        the operators torch has, otherwise through host copies of the slices."""
        import torch
        fns = {N.SUM: torch.add, N.MAX: torch.maximum, N.MIN: torch.minimum, N.PROD: torch.mul}
        acc = arrs[dst][f:t]
        order = [i for i in range(len(arrs)) if i != dst]
        fn = fns.get(operator.code) if operator.is_native else None
        if fn is not None:
            for i in order:
                fn(acc, arrs[i][f:t], out=acc)
            return
        hs = [arrs[dst][f:t].cpu()] + [arrs[i][f:t].cpu() for i in order]
        self._local_reduce(hs, operand, operator, 0, t - f, 0)
        acc.copy_(hs[0])

    def _local_copy(self, dst, src, f, t, operand):
        if f >= t or dst is src:
            return
        if _is_torch(dst):
            dst[f:t].copy_(src[f:t])
        else:
            d, s = self._host(dst, operand), self._host(src, operand)
            d[f:t] = s[f:t]

    # ------------------------------------------------------------------ array collectives
    def gatherArray(self, arrData, operand, sendfroms, sendtos, rootRank, rootThreadId):
        """2-D ranges [slaveNum][threadNum]; result lands in the root thread's array (TCS:448)."""
        if self._T == 1:
            return self.processCommSlave.gatherArray(arrData, operand, CommUtils.getProcessFroms(sendfroms),
                                                     CommUtils.getProcessTos(sendtos), rootRank)
        CommUtils.isfromsTosLegal(sendfroms, sendtos, self._T, size=len(arrData))
        rr = self._real_root(rootRank, rootThreadId)
        r = self.getRank()
        if self._gpu_segments("thread_gather", arrData, operand, None,
                              [(sendfroms[q][t], sendtos[q][t], q, t, rootRank, rootThreadId)
                               for q in range(self.getSlaveNum()) for t in range(self._T)], rr):
            return arrData
        if self._native_thread(_GATHER, arrData, operand, None, lo=sendfroms, hi=sendtos, rootRank=rootRank,
                               rootThreadId=rootThreadId):
            return arrData

        def run(arrs):
            for t in range(self._T):  # thread merge into the delegate's array (DOP:400-419)
                self._local_copy(arrs[rr], arrs[t], sendfroms[r][t], sendtos[r][t], operand)
            self.processCommSlave.gatherArray(arrs[rr], operand, CommUtils.getProcessFroms(sendfroms),
                                              CommUtils.getProcessTos(sendtos), rootRank)
            return None
        self._delegate(arrData, run, rr)
        return arrData

    def allgatherArray(self, arrData, operand, sendfroms, sendtos):
        if self._T == 1:
            return self.processCommSlave.allgatherArray(arrData, operand, CommUtils.getProcessFroms(sendfroms),
                                                        CommUtils.getProcessTos(sendtos))
        CommUtils.isfromsTosLegal(sendfroms, sendtos, self._T, size=len(arrData))
        r = self.getRank()
        pf, pt = CommUtils.getProcessFroms(sendfroms), CommUtils.getProcessTos(sendtos)
        if self._gpu_segments("thread_allgather", arrData, operand, None,
                              [(sendfroms[q][t], sendtos[q][t], q, t, -1, -1)
                               for q in range(self.getSlaveNum()) for t in range(self._T)]):
            return arrData
        if self._native_thread(_ALL_GATHER, arrData, operand, None, lo=sendfroms, hi=sendtos):
            return arrData

        def run(arrs):
            for t in range(1, self._T):
                self._local_copy(arrs[0], arrs[t], sendfroms[r][t], sendtos[r][t], operand)
            self.processCommSlave.allgatherArray(arrs[0], operand, pf, pt)
            for t in range(1, self._T):  # full copy-out to every thread (TCS:676, DOP:386-393)
                for q in range(self.getSlaveNum()):
                    self._local_copy(arrs[t], arrs[0], pf[q], pt[q], operand)
            return None
        self._delegate(arrData, run, 0)
        return arrData

    def broadcastArray(self, arrData, operand, from_, to, rootRank, rootThreadId):
        if self._T == 1:
            return self.processCommSlave.broadcastArray(arrData, operand, from_, to, rootRank)
        rr = self._real_root(rootRank, rootThreadId)
        CommUtils.isFromToLegal(from_, to, len(arrData))
        if self._gpu_segments("thread_broadcast", arrData, operand, None,
                              [(from_, to, rootRank, rootThreadId, -1, -1)], rr):
            return arrData
        if self._native_thread(_BROADCAST, arrData, operand, None, from_, to, rootRank=rootRank,
                               rootThreadId=rootThreadId):
            return arrData

        def run(arrs):
            self.processCommSlave.broadcastArray(arrs[rr], operand, from_, to, rootRank)
            for t in range(self._T):
                self._local_copy(arrs[t], arrs[rr], from_, to, operand)
            return None
        self._delegate(arrData, run, rr)
        return arrData

    def scatterArray(self, arrData, operand, recvfroms, recvtos, rootRank, rootThreadId):
        if self._T == 1:
            return self.processCommSlave.scatterArray(arrData, operand, CommUtils.getProcessFroms(recvfroms),
                                                      CommUtils.getProcessTos(recvtos), rootRank)
        CommUtils.isfromsTosLegal(recvfroms, recvtos, self._T, size=len(arrData))
        rr = self._real_root(rootRank, rootThreadId)
        r = self.getRank()
        if self._gpu_segments("thread_scatter", arrData, operand, None,
                              [(recvfroms[q][t], recvtos[q][t], rootRank, rootThreadId, q, t)
                               for q in range(self.getSlaveNum()) for t in range(self._T)], rr):
            return arrData
        if self._native_thread(_SCATTER, arrData, operand, None, lo=recvfroms, hi=recvtos, rootRank=rootRank,
                               rootThreadId=rootThreadId):
            return arrData

        def run(arrs):
            self.processCommSlave.scatterArray(arrs[rr], operand, CommUtils.getProcessFroms(recvfroms),
                                               CommUtils.getProcessTos(recvtos), rootRank)
            for t in range(self._T):
                self._local_copy(arrs[t], arrs[rr], recvfroms[r][t], recvtos[r][t], operand)
            return None
        self._delegate(arrData, run, rr)
        return arrData

    def reduceScatterArray(self, arrData, operand, operator, from_, counts):
        """counts is [slaveNum][threadNum]; thread (r,t) ends with its fully reduced slice (TCS:1174)."""
        if self._T == 1:
            return self.processCommSlave.reduceScatterArray(arrData, operand, operator, from_,
                                                            [c[0] for c in counts])
        CommUtils.isFromCountsLegal(from_, counts, len(arrData))
        r = self.getRank()
        pcounts = [sum(c) for c in counts]
        total_to = from_ + sum(pcounts)
        if _is_cuda(arrData):
            segs, off = [], from_
            for q in range(self.getSlaveNum()):
                for t in range(self._T):
                    segs.append((off, off + counts[q][t], -1, -1, q, t))
                    off += counts[q][t]
            if self._gpu_segments("thread_reduce_scatter", arrData, operand, operator, segs):
                return arrData
        if self._native_thread(_REDUCE_SCATTER, arrData, operand, operator, from_, total_to, counts=counts):
            return arrData

        def run(arrs):
            self._local_reduce(arrs, operand, operator, from_, total_to, 0)
            self.processCommSlave.reduceScatterArray(arrs[0], operand, operator, from_, pcounts)
            off = from_ + sum(pcounts[:r])
            for t in range(self._T):
                self._local_copy(arrs[t], arrs[0], off, off + counts[r][t], operand)
                off += counts[r][t]
            return None
        self._delegate(arrData, run, 0)
        return arrData

    def reduceArray(self, arrData, operand, operator, from_, to, rootRank, rootThreadId):
        """The root thread's array gets the reduction over all slaveNum x threadNum arrays (TCS:1396). As in the
        reference (reduce-scatter + gather in place, PCS:1390-1416) [from, to) of the other arrays is scratch: the
        host path reduces into the delegate thread's array, the GPU relay (beyond 2 ranks and 4 MiB) rests partial
        results in thread 0 of every non-root rank."""
        if self._T == 1:
            return self.processCommSlave.reduceArray(arrData, operand, operator, from_, to, rootRank)
        rr = self._real_root(rootRank, rootThreadId)
        CommUtils.isFromToLegal(from_, to, len(arrData))
        if self._gpu_segments("thread_reduce", arrData, operand, operator,
                              [(from_, to, -1, -1, rootRank, rootThreadId)], rr):
            return arrData
        if self._native_thread(_REDUCE, arrData, operand, operator, from_, to, rootRank=rootRank,
                               rootThreadId=rootThreadId):
            return arrData

        def run(arrs):
            self._local_reduce(arrs, operand, operator, from_, to, rr)
            self.processCommSlave.reduceArray(arrs[rr], operand, operator, from_, to, rootRank)
            return None
        self._delegate(arrData, run, rr)
        return arrData

    def allreduceArray(self, arrData, operand, operator, from_=0, to=None, resultThread=None):
        """All slaveNum x threadNum arrays are reduced; every thread's array holds the result
        (TCS:1915-1963). CUDA tensors take the fused thread-reduce + inter-GPU + fan-out kernel; pinned
        host tensors are streamed through the GPU plane when it is initialised.

        resultThread (an addition; None = the reference's behaviour): when the T threads of a process
        share the result anyway, name the one thread whose array must hold it on every rank; the other
        threads' arrays are then unspecified (input or result). On the host-array pipeline this cuts the
        device->host traffic from T copies of the result to one."""
        if to is None:
            to = len(arrData)
        if resultThread is not None and not 0 <= resultThread < self._T:
            raise Mp4jException("resultThread out of range")
        if self._T == 1:
            return self.processCommSlave.allreduceArray(arrData, operand, operator, from_, to)
        CommUtils.isFromToLegal(from_, to, len(arrData))
        if self.processCommSlave._gpu_host_ok(arrData, operand, operator):
            # pinned host tensors with the GPU plane up: thread 0 streams all T arrays through the GPU
            # (H2D || fused kernel || D2H pipeline in the library); every thread's array gets the result
            self.gpu.thread_native_allreduce(self._tg, self.getThreadId(), arrData, operand, operator, from_, to,
                                             host=True, fanout=resultThread is None, dst=resultThread or 0)
            return arrData
        if not _is_cuda(arrData):
            # host arrays + predefined operator: native path in which ALL T threads take part (each
            # reduces its 1/T slice, each copies its own result) instead of one delegate doing everything
            a = self.processCommSlave._native_host(arrData, operand, operator)
            if a is not None:
                check(N.lib.mp4j_thread_allreduce(self._tg, self.processCommSlave._h, self.getThreadId(),
                                                  a.ctypes.data, operand.code, operator.code, from_, to))
                return arrData

        if _is_cuda(arrData):
            # CUDA tensors: the fused thread-reduce + inter-GPU + fan-out kernel, launched by thread 0 from the
            # native rendezvous (every thread calls in with its own tensor). Without a GPU plane (ranks on several
            # machines) the generic composition below runs with the process step staged through the host plane.
            plane = self._cuda_plane()
            if plane is not None:
                plane.thread_native_allreduce(self._tg, self.getThreadId(), arrData, operand, operator, from_, to)
                return arrData

        def run(arrs):
            self._local_reduce(arrs, operand, operator, from_, to, 0)
            self.processCommSlave.allreduceArray(arrs[0], operand, operator, from_, to)
            for t in range(1, self._T):
                self._local_copy(arrs[t], arrs[0], from_, to, operand)
            return None
        self._delegate(arrData, run, 0)
        return arrData

    def allreduceArrayRpc(self, arrData, operand, operator):
        if self._T == 1:
            return self.processCommSlave.allreduceArrayRpc(arrData, operand, operator)
        n = len(arrData)

        def run(arrs):
            self._local_reduce(arrs, operand, operator, 0, n, 0)
            self.processCommSlave.allreduceArrayRpc(arrs[0], operand, operator)
            for t in range(1, self._T):
                self._local_copy(arrs[t], arrs[0], 0, n, operand)
            return None
        self._delegate(arrData, run, 0)
        return arrData

    # ------------------------------------------------------------------ scalars
    def broadcast(self, value, operand, rootRank, rootThreadId):
        if self._T == 1:
            return self.processCommSlave.broadcast(value, operand, rootRank)
        rr = self._real_root(rootRank, rootThreadId)
        return self._delegate(value, lambda vals: self.processCommSlave.broadcast(vals[rr], operand, rootRank), rr)

    def _thread_fold(self, vals, operator, first):
        order = [first] + [i for i in range(len(vals)) if i != first]
        acc = vals[order[0]]
        for i in order[1:]:
            acc = operator.apply(acc, vals[i])
        return acc

    def reduce(self, value, operand, operator, rootRank, rootThreadId):
        """Root thread gets the reduction; every other thread gets its own input back (TCS:1433)."""
        if self._T == 1:
            return self.processCommSlave.reduce(value, operand, operator, rootRank)
        rr = self._real_root(rootRank, rootThreadId)
        is_root_proc = self.getRank() == rootRank

        def run(vals):
            acc = operand.cast(self._thread_fold(vals, operator, rr))
            return self.processCommSlave.reduce(acc, operand, operator, rootRank)
        res = self._delegate(value, run, rr)
        return res if (is_root_proc and self.getThreadId() == rr) else value

    def allreduce(self, value, operand, operator):
        if self._T == 1:
            return self.processCommSlave.allreduce(value, operand, operator)
        return self._delegate(value, lambda vals: self.processCommSlave.allreduce(
            operand.cast(self._thread_fold(vals, operator, 0)), operand, operator), 0)

    def allreduceRpc(self, value, operand, operator):
        if self._T == 1:
            return self.processCommSlave.allreduceRpc(value, operand, operator)
        return self._delegate(value, lambda vals: self.processCommSlave.allreduceRpc(
            operand.cast(self._thread_fold(vals, operator, 0)), operand, operator), 0)

    # ------------------------------------------------------------------ maps
    @staticmethod
    def _merge_maps(maps, operator=None):
        out = {}
        if operator is None:
            for m in maps:
                out.update(m)
            return out
        ap = operator.apply
        for m in maps:
            for k, v in m.items():
                out[k] = ap(out[k], v) if k in out else v
        return out

    def gatherMap(self, mapData, operand, rootRank, rootThreadId):
        if self._T == 1:
            return self.processCommSlave.gatherMap(mapData, operand, rootRank)
        rr = self._real_root(rootRank, rootThreadId)
        is_root_proc = self.getRank() == rootRank
        res = self._delegate(mapData, lambda maps: self.processCommSlave.gatherMap(
            self._merge_maps(maps), operand, rootRank), rr)
        return res if (is_root_proc and self.getThreadId() == rr) else None

    def allgatherMap(self, mapData, operand):
        if self._T == 1:
            return self.processCommSlave.allgatherMap(mapData, operand)
        return self._delegate(mapData, lambda maps: self.processCommSlave.allgatherMap(
            self._merge_maps(maps), operand), 0)

    def broadcastMap(self, mapData, operand, rootRank, rootThreadId):
        if self._T == 1:
            return self.processCommSlave.broadcastMap(mapData, operand, rootRank)
        rr = self._real_root(rootRank, rootThreadId)
        return self._delegate(mapData, lambda maps: self.processCommSlave.broadcastMap(
            maps[rr], operand, rootRank), rr)

    def scatterMap(self, mapDataListList, operand, rootRank, rootThreadId):
        """Root thread holds [slaveNum][threadNum] maps; thread (r,t) receives map [r][t] (TCS:1097)."""
        if self._T == 1:
            lst = [m[0] for m in mapDataListList] if mapDataListList is not None else None
            return self.processCommSlave.scatterMap(lst, operand, rootRank)
        rr = self._real_root(rootRank, rootThreadId)
        res = self._delegate(mapDataListList, lambda lls: self.processCommSlave._scatterMapSpecial(
            lls[rr], operand, rootRank), rr)
        return res[self.getThreadId()]

    def reduceScatterMap(self, mapDataListList, operand, operator):
        """Every thread contributes [slaveNum][threadNum] maps; thread (r,t) gets reduced map [r][t]."""
        p = self.getSlaveNum()
        if self._T == 1:
            return self.processCommSlave.reduceScatterMap([m[0] for m in mapDataListList], operand, operator)

        def run(lls):
            merged = [[self._merge_maps([lls[s][q][t] for s in range(self._T)], operator)
                       for t in range(self._T)] for q in range(p)]
            if p == 1:
                return merged[0]
            mm = ProcessCommSlave._merge_map(operator)

            def merge(local, incoming):
                return [mm(a, b) for a, b in zip(local, incoming)]
            return self.processCommSlave._ring_reduce_scatter(merged, merge, operand)
        res = self._delegate(mapDataListList, run, 0)
        return res[self.getThreadId()]

    def reduceMap(self, mapData, operand, operator, rootRank, rootThreadId):
        if self._T == 1:
            return self.processCommSlave.reduceMap(mapData, operand, operator, rootRank)
        rr = self._real_root(rootRank, rootThreadId)
        is_root_proc = self.getRank() == rootRank
        res = self._delegate(mapData, lambda maps: self.processCommSlave.reduceMap(
            self._merge_maps(maps, operator), operand, operator, rootRank), rr)
        return res if (is_root_proc and self.getThreadId() == rr) else None  # TCS:1583

    def allreduceMap(self, mapData, operand, operator):
        """Result map is SHARED by the threads of a process - clone before mutating (README:254)."""
        if self._T == 1:
            return self.processCommSlave.allreduceMap(mapData, operand, operator)
        pcs = self.processCommSlave
        if pcs._gpu_map_ok(operand, operator):
            # GPU path: the T thread maps go to the device as they are - the merge kernel folds equal keys of
            # different threads exactly like equal keys of different ranks (TCS:2179-2224 without the host merge)
            def run_gpu(maps):
                out = pcs._gpu_allreduce_map(maps, operand, operator)
                return out if out is not None else pcs.allreduceMap(self._merge_maps(maps, operator), operand, operator)
            return self._delegate(mapData, run_gpu, 0)
        return self._delegate(mapData, lambda maps: self.processCommSlave.allreduceMap(
            self._merge_maps(maps, operator), operand, operator), 0)

    # ------------------------------------------------------------------ set / list helpers (TCS:1607-1863, 2241-2455)
    def reduceMapSetUnion(self, mapData, rootRank, rootThreadId, elementSerializer=None, elementType=None):
        return self.reduceMap(mapData, ProcessCommSlave._obj_operand(elementSerializer, elementType),
                              ProcessCommSlave._UNION, rootRank, rootThreadId)

    def reduceSetUnion(self, setData, rootRank, rootThreadId, elementSerializer=None, elementType=None):
        m = self.reduceMapSetUnion({"key": setData}, rootRank, rootThreadId, elementSerializer, elementType)
        return None if m is None else m.get("key")

    def reduceMapSetIntersection(self, mapData, rootRank, rootThreadId, elementSerializer=None, elementType=None):
        return self.reduceMap(mapData, ProcessCommSlave._obj_operand(elementSerializer, elementType),
                              ProcessCommSlave._INTERSECT, rootRank, rootThreadId)

    def reduceSetIntersection(self, setData, rootRank, rootThreadId, elementSerializer=None, elementType=None):
        m = self.reduceMapSetIntersection({"key": setData}, rootRank, rootThreadId, elementSerializer, elementType)
        return None if m is None else m.get("key")

    def reduceMapListConcat(self, mapData, rootRank, rootThreadId, elementSerializer=None, elementType=None):
        return self.reduceMap(mapData, ProcessCommSlave._obj_operand(elementSerializer, elementType),
                              ProcessCommSlave._CONCAT, rootRank, rootThreadId)

    def reduceListConcat(self, listData, rootRank, rootThreadId, elementSerializer=None, elementType=None):
        m = self.reduceMapListConcat({"key": listData}, rootRank, rootThreadId, elementSerializer, elementType)
        return None if m is None else m.get("key")

    def allreduceMapSetUnion(self, mapData, elementSerializer=None, elementType=None):
        return self.allreduceMap(mapData, ProcessCommSlave._obj_operand(elementSerializer, elementType),
                                 ProcessCommSlave._UNION)

    def allreduceSetUnion(self, setData, elementSerializer=None, elementType=None):
        return self.allreduceMapSetUnion({"key": setData}, elementSerializer, elementType)["key"]

    def allreduceMapSetIntersection(self, mapData, elementSerializer=None, elementType=None):
        return self.allreduceMap(mapData, ProcessCommSlave._obj_operand(elementSerializer, elementType),
                                 ProcessCommSlave._INTERSECT)

    def allreduceSetIntersection(self, setData, elementSerializer=None, elementType=None):
        return self.allreduceMapSetIntersection({"key": setData}, elementSerializer, elementType)["key"]

    def allreduceMapListConcat(self, mapData, elementSerializer=None, elementType=None):
        return self.allreduceMap(mapData, ProcessCommSlave._obj_operand(elementSerializer, elementType),
                                 ProcessCommSlave._CONCAT)

    def allreduceListConcat(self, listData, elementSerializer=None, elementType=None):
        return self.allreduceMapListConcat({"key": listData}, elementSerializer, elementType)["key"]


def _install_process_passthroughs():
    """xxxProcess(...) = ProcessCommSlave.xxx(...) for 'only one thread per process calls this'
    (TCS:429, 526, 606, ... 2438)."""
    names = ["gatherArray", "gatherMap", "allgatherArray", "allgatherMap", "broadcastArray", "broadcast",
             "broadcastMap", "scatterArray", "scatterMap", "reduceScatterArray", "reduceScatterMap",
             "reduceArray", "reduce", "reduceMap", "allreduceArray", "allreduceArrayRpc", "allreduce",
             "allreduceRpc", "allreduceMap",
             "reduceMapSetUnion", "reduceSetUnion", "reduceMapSetIntersection", "reduceSetIntersection",
             "reduceMapListConcat", "reduceListConcat", "allreduceMapSetUnion", "allreduceSetUnion",
             "allreduceMapSetIntersection", "allreduceSetIntersection", "allreduceMapListConcat",
             "allreduceListConcat"]
    for name in names:
        def make(nm):
            def passthrough(self, *a, **kw):
                return getattr(self.processCommSlave, nm)(*a, **kw)
            passthrough.__name__ = nm + "Process"
            passthrough.__doc__ = "Pass-through to ProcessCommSlave.%s (one thread per process calls)." % nm
            return passthrough
        setattr(ThreadCommSlave, name + "Process", make(name))


_install_process_passthroughs()
