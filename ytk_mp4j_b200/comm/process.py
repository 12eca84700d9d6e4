"""ProcessCommSlave: the process-level communicator (parity: J/comm/ProcessCommSlave.java).

Three data paths behind one API:
  * CUDA tensors  -> GPU data plane (csrc/gpu): hand-written sm_100a kernels over peer memory;
  * numpy / CPU tensors with a predefined operator -> C++ host collectives over TCP
    (csrc/host/process_comm.cc: ring reduce-scatter / ring allgather / binomial gather+scatter);
  * Python lists (String/Object operands), maps and custom Python operators -> the same four
    algorithms written here over the C++ byte transport, with the operand's serializer
    (pickle by default, optional Deflate) as the codec.
`slaveNum == 1` short-circuits every collective to "return the input" (PCS:545,630,751,1179,1734).
"""
import ctypes as C
import os
import socket
import threading
import time
import traceback

import numpy as np

from .. import _native as N
from ..exception import Mp4jException, check
from ..operands import Operand, Operands
from ..operators import Operator, custom
from ..utils.comm_utils import CommUtils

import sys


def _torch():
    """torch is optional for the host paths and slow to import: only look at it once the caller
    has imported it (a tensor argument implies that)."""
    return sys.modules.get("torch")


def _is_torch(x):
    t = _torch()
    return t is not None and isinstance(x, t.Tensor)


def _is_cuda(x):
    return _is_torch(x) and x.is_cuda


def _is_pinned_host(x):
    """A pinned CPU tensor: the caller's way of asking for the GPU plane's host-array pipeline."""
    return _is_torch(x) and not x.is_cuda and x.dim() == 1 and x.is_contiguous() and x.is_pinned()


def java_string_hash(key):
    """java.lang.String.hashCode for str keys; ints hash to themselves (Integer.hashCode)."""
    if isinstance(key, str):
        b = key.encode("utf-8")
        return N.lib.mp4j_java_string_hash(b, len(b))
    if isinstance(key, (int, np.integer)):
        k = int(key) & 0xFFFFFFFFFFFFFFFF
        h = (k ^ (k >> 32)) & 0xFFFFFFFF  # Long.hashCode
        return h - (1 << 32) if h & 0x80000000 else h
    return hash(key) & 0x7FFFFFFF


def key_partition(key, p):
    """Owner rank of a map key: hashCode % p with negatives wrapped (PCS:2066-2070)."""
    h = java_string_hash(key)
    m = int(np.fmod(h, p))  # Java remainder keeps the sign of the dividend
    return m + p if m < 0 else m


class _MapCodec:
    """Wire format of Map<String, primitive> payloads on the native blob collectives: [tag 0][u64 n][u64 key bytes]
    [NUL-separated UTF-8 keys][n values] - no per-entry objects (the role of map_host.cc's Part for the collectives
    that only MOVE maps: gatherMap, allgatherMap, broadcastMap, scatterMap). Anything that is not a plain dict of
    str keys (or has NUL inside a key) falls back to the operand's own codec behind tag 1."""

    def __init__(self, operand):
        self.operand = operand
        self.dtype = np.dtype(operand.np_dtype)

    def dumps(self, m):
        if isinstance(m, dict):
            n = len(m)
            try:
                keys = "\0".join(m.keys()).encode("utf-8")
            except TypeError:
                keys = None
            if keys is not None and (n == 0 or keys.count(b"\0") == n - 1):
                vals = np.fromiter(m.values(), dtype=self.dtype, count=n).tobytes()
                return b"\0" + n.to_bytes(8, "little") + len(keys).to_bytes(8, "little") + keys + vals
        return b"\1" + self.operand.dumps(m)

    def loads(self, raw):
        if raw[:1] == b"\1":
            return self.operand.loads(raw[1:])
        n = int.from_bytes(raw[1:9], "little")
        kl = int.from_bytes(raw[9:17], "little")
        if n == 0:
            return {}
        keys = raw[17:17 + kl].decode("utf-8").split("\0")
        vals = np.frombuffer(raw, dtype=self.dtype, count=n, offset=17 + kl).tolist()
        return dict(zip(keys, vals))


class ProcessCommSlave:
    def __init__(self, loginName, masterHost, masterPort, rank_hint=None, host=None):
        """Connects to the master, performs the rendezvous and learns rank / peers (PCS:143-210).
        rank_hint (or env MP4J_RANK): ask the master for a specific rank; honoured when every
        slave passes a distinct valid hint, else the reference rule applies (rank = position of
        "host###port" in the sorted address list)."""
        if rank_hint is None and os.environ.get("MP4J_RANK"):
            rank_hint = int(os.environ["MP4J_RANK"])
        self._h = C.c_void_p()
        check(N.lib.mp4j_comm_create((loginName or "").encode(), masterHost.encode(), int(masterPort),
                                     -1 if rank_hint is None else int(rank_hint),
                                     host.encode() if host else None, C.byref(self._h)))
        self._rank = N.lib.mp4j_comm_rank(self._h)
        self._n = N.lib.mp4j_comm_size(self._h)
        self._prefix = "[rank=%d] " % self._rank  # PCS:166
        self._closed = False
        self._gpu = None
        self._lock = threading.RLock()
        self.stats = []  # per-collective records from the GPU plane
        self._plane_blocker = None  # '' = the GPU plane can span this job, text = why not (decided on first use)
        self.info("this slave init finished!", onlyRank0=False)  # PCS:209

    # ------------------------------------------------------------------ identity / lifecycle
    def getRank(self):
        return self._rank

    def getSlaveNum(self):
        return self._n

    rank = property(getRank)
    size = property(getSlaveNum)

    @property
    def host_transport(self):
        """"shm" (ranks share a machine: host arrays move through shared-memory windows), "tcp", or "tcp+shm"
        (several machines with several ranks each: allreduce is hierarchical - shared memory inside a machine,
        TCP between the machines)."""
        return N.lib.mp4j_comm_host_transport(self._h).decode()

    def barrier(self):
        check(N.lib.mp4j_comm_barrier(self._h))

    def _log(self, level, msg):
        check(N.lib.mp4j_comm_log(self._h, level, msg.encode("utf-8", "replace")))

    def info(self, info, onlyRank0=True):
        """Forwards to the master's log sink; by default only rank 0 sends (PCS:289-310)."""
        if onlyRank0 and self._rank != 0:
            return
        self._log(N.LOG_INFO, self._prefix + str(info))

    def debug(self, debug, onlyRank0=True):
        if onlyRank0 and self._rank != 0:
            return
        self._log(N.LOG_DEBUG, self._prefix + str(debug))

    def error(self, error):
        self._log(N.LOG_ERROR, self._prefix + str(error))

    def exception(self, e, sleep_s=5.0):
        """Reports the stack trace to the master, waits, then close(1) (PCS:360-373)."""
        trace = "".join(traceback.format_exception(type(e), e, e.__traceback__))
        try:
            self.error("slave exception:" + trace)
        finally:
            time.sleep(sleep_s)
            self.close(1)

    def writeFile(self, content, fileName):
        """Creates the file ON THE MASTER (PCS:279, SRV:302-316)."""
        check(N.lib.mp4j_comm_write_file(self._h, str(content).encode(), str(fileName).encode()))

    def close(self, code=0):
        with self._lock:
            if self._closed:
                return
            self._closed = True
            if self._gpu is not None:
                try:
                    self._gpu.close()
                finally:
                    self._gpu = None
            check(N.lib.mp4j_comm_close(self._h, int(code)))

    def isClosed(self):
        return self._closed

    def __del__(self):
        try:
            if self._h:
                if self._gpu is not None:
                    self._gpu.close()
                N.lib.mp4j_comm_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    # ------------------------------------------------------------------ GPU plane
    def init_gpu(self, device=None, heap_bytes=None, staging_bytes=None):
        """Sets up the GPU data plane: symmetric heap, peer mapping, signal pads, multicast.
        Collective: every rank must call it. Called lazily by the first CUDA-tensor collective.
        The plane is one NVLink domain: when the ranks of the job sit on several machines (or MP4J_GPU_PLANE=0)
        it cannot be built and this raises; collectives on CUDA tensors then go through the host plane
        (see _staged)."""
        if self._gpu is None:
            why = self._gpu_plane_blocker()
            if why:
                raise Mp4jException("the GPU plane cannot be set up: " + why)
            from ..parallel.gpu_plane import GpuPlane
            self._gpu = GpuPlane(self, device=device, heap_bytes=heap_bytes, staging_bytes=staging_bytes)
        return self._gpu

    def _gather_all_bytes(self, data):
        """Every rank's byte string, in rank order, on every rank (through the master mailbox)."""
        ptr, n = C.c_void_p(), C.c_size_t()
        check(N.lib.mp4j_comm_gather_all(self._h, data, len(data), C.byref(ptr), C.byref(n)))
        raw, out, off = N.take_bytes(ptr, n), [], 0
        for _ in range(self._n):
            ln = int.from_bytes(raw[off:off + 8], "little")
            out.append(raw[off + 8:off + 8 + ln])
            off += 8 + ln
        return out

    def _gpu_plane_blocker(self):
        """Why the GPU plane cannot span this job ('' = it can). Decided once, by all ranks together (collective):
        the plane maps every rank's heap into every other rank (VMM handles passed as file descriptors, NVLink
        peer access), which works inside one machine only."""
        if self._plane_blocker is None:
            try:
                with open("/proc/sys/kernel/random/boot_id") as f:
                    boot = f.read().strip()
            except OSError:
                boot = ""
            # MP4J_NODE_ID (tests: several "machines" on one box) is part of the machine identity, as in the host plane
            me = "%s|%s#%s|%s" % (socket.gethostname(), boot, os.environ.get("MP4J_NODE_ID", ""),
                                  os.environ.get("MP4J_GPU_PLANE", "1"))
            everyone = [b.decode() for b in self._gather_all_bytes(me.encode())] if self._n > 1 else [me]
            machines = {e.rsplit("|", 1)[0] for e in everyone}
            if any(e.rsplit("|", 1)[1] == "0" for e in everyone):
                self._plane_blocker = "MP4J_GPU_PLANE=0"
            elif self._n > N.MAX_RANKS:
                self._plane_blocker = "the job has %d ranks and the plane spans at most %d (one NVLink domain)" % (
                    self._n, N.MAX_RANKS)
            elif len(machines) > 1:
                self._plane_blocker = "the %d ranks run on %d machines and the plane is one NVLink domain" % (
                    self._n, len(machines))
            else:
                self._plane_blocker = ""
        return self._plane_blocker

    def _cuda_plane(self):
        """The GPU plane for a collective on CUDA tensors, or None when it cannot exist for this job."""
        if self._gpu is not None:
            return self._gpu
        return None if self._gpu_plane_blocker() else self.init_gpu()

    def _staged(self, t, fn, operator=None):
        """Collective on a CUDA tensor without a GPU plane (ranks on several machines): device -> host copy, the
        host-plane collective (shared memory inside a machine, TCP between machines), host -> device copy. The
        multi-machine form of every array collective; a hierarchical form (NVLink kernels inside a machine, TCP
        between machines) is the faster design this falls short of."""
        if operator is not None and getattr(operator, "expr", None) and operator.code is None:
            raise Mp4jException("device expressions need the GPU plane; use a predefined or a Python operator here")
        h = t.detach().to("cpu")
        fn(h)
        t.copy_(h)
        return t

    @property
    def gpu(self):
        return self.init_gpu()

    def empty(self, n, operand):
        """Allocates an n-element tensor inside the symmetric heap: collectives on it are
        zero-copy (no staging)."""
        return self.gpu.empty(n, operand)

    # ------------------------------------------------------------------ helpers
    def _host_view(self, arr, operand):
        """Returns (numpy view sharing memory, original) for host arrays of primitive operands."""
        if _is_torch(arr):
            a = arr.detach().numpy() if arr.dtype != _torch().bfloat16 else arr.detach().view(_torch().int16).numpy()
        elif isinstance(arr, np.ndarray):
            a = arr
        else:
            return None
        if a.ndim != 1 or not a.flags["C_CONTIGUOUS"]:
            raise Mp4jException("array collectives need 1-D C-contiguous arrays")
        if operand.np_dtype is not None and a.dtype != np.dtype(operand.np_dtype):
            raise Mp4jException("array dtype %s does not match operand %s" % (a.dtype, operand.name))
        return a

    def _gpu_host_ok(self, arr, operand, operator):
        """Pinned host tensor + initialised GPU plane + an (operand, operator) pair the fused kernels
        know: the array is streamed through the GPU instead of the host plane."""
        return (self._gpu is not None and _is_pinned_host(arr) and operand.is_primitive and operator is not None
                and operator.is_native and operand.code in (N.F64, N.F32, N.I64, N.I32))

    def _native_host(self, arr, operand, operator):
        if not operand.is_primitive:
            return None
        if operator is not None and not operator.is_native:
            return None
        return self._host_view(arr, operand)

    @staticmethod
    def _merge_seq(operator, operand):
        """Element-wise merge op(local, incoming) for list / numpy segments (DOP:192-198)."""
        def merge(local, incoming):
            if isinstance(local, np.ndarray):
                if operator.is_native:
                    out = np.array(local, copy=True)
                    inc = np.ascontiguousarray(incoming, dtype=out.dtype)
                    check(N.lib.mp4j_host_reduce_into(operand.code, operator.code, out.ctypes.data,
                                                      inc.ctypes.data, out.size))
                    return out
                if operator.vectorized:
                    return np.asarray(operator.fn(local, incoming), dtype=local.dtype)
                return np.array([operator.fn(a, b) for a, b in zip(local.tolist(), incoming.tolist())],
                                dtype=local.dtype)
            return [operator.apply(a, b) for a, b in zip(local, incoming)]
        return merge

    @staticmethod
    def _merge_map(operator):
        def merge(local, incoming):
            ap = operator.apply
            for k, v in incoming.items():
                local[k] = ap(local[k], v) if k in local else v
            return local
        return merge

    # ------------------------------------------------------------------ generic transport (bytes)
    def _send(self, dest, payload, operand):
        data = operand.dumps(payload)
        check(N.lib.mp4j_comm_send_bytes(self._h, dest, data, len(data)))

    def _recv(self, src, operand):
        ptr, n = C.c_void_p(), C.c_size_t()
        check(N.lib.mp4j_comm_recv_bytes(self._h, src, C.byref(ptr), C.byref(n)))
        return operand.loads(N.take_bytes(ptr, n))

    def _sendrecv(self, dest, payload, src, operand):
        data = operand.dumps(payload)
        ptr, n = C.c_void_p(), C.c_size_t()
        check(N.lib.mp4j_comm_sendrecv_bytes(self._h, dest, data, len(data), src, C.byref(ptr), C.byref(n)))
        return operand.loads(N.take_bytes(ptr, n))

    # the four data-moving algorithms of the reference on opaque payloads: the ring / tree runs in C++
    # (csrc/host/blob_coll.cc); this side only serializes, deserializes and merges (the operand's job in the
    # reference too, ObjectOperand.java:114-156). Inner tree nodes forward peers' payloads as raw bytes.
    def _blob_collective(self, kind, root, get_part, put_part):
        """get_part(i) -> payload object of local part i; put_part(i, payload) consumes a received payload."""
        hold, err = [None], []

        def c_get(_user, part, data, n):
            try:
                b = get_part(part)
                hold[0] = b  # keeps the bytes alive until the next callback
                data[0] = C.cast(C.c_char_p(b), C.c_void_p)
                n[0] = len(b)
                return 0
            except BaseException as e:  # noqa: BLE001 - reported after the native call returns
                err.append(e)
                return 1

        def c_put(_user, part, data, n):
            try:
                put_part(part, C.string_at(data, n))
                return 0
            except BaseException as e:  # noqa: BLE001
                err.append(e)
                return 1
        g, pcb = N.BLOB_GET(c_get), N.BLOB_PUT(c_put)
        rc = N.lib.mp4j_host_blob_collective(self._h, kind, root, g, pcb, None)
        if err:
            raise err[0] if isinstance(err[0], Mp4jException) else Mp4jException(repr(err[0]))
        check(rc)

    def _ring_reduce_scatter(self, parts, merge, operand):
        """parts[i] = this rank's contribution to rank i's result (PCS:1329-1373)."""
        def put(i, raw):
            parts[i] = merge(parts[i], operand.loads(raw))
        self._blob_collective(N.REDUCE_SCATTER, 0, lambda i: operand.dumps(parts[i]), put)
        return parts[self._rank]

    def _ring_allgather(self, mine, operand):
        """Returns the p payloads indexed by origin rank (PCS:694-737)."""
        out = [None] * self._n
        out[self._rank] = mine

        def put(i, raw):
            out[i] = operand.loads(raw)
        self._blob_collective(N.ALL_GATHER, 0, lambda i: operand.dumps(out[i]), put)
        return out

    def _tree_gather(self, mine, root, operand):
        """Binomial gather; returns {rank: payload} (complete on root)."""
        acc = {self._rank: mine}

        def put(i, raw):
            acc[i] = operand.loads(raw)
        self._blob_collective(N.GATHER, root, lambda i: operand.dumps(mine), put)
        return acc

    def _tree_scatter(self, payloads, root, operand):
        """Binomial scatter of payloads[0..p) held by root; returns this rank's payload."""
        box = [payloads[self._rank] if self._rank == root else None]

        def put(i, raw):
            box[0] = operand.loads(raw)
        self._blob_collective(N.SCATTER, root, lambda i: operand.dumps(payloads[i]), put)
        return box[0]

    @staticmethod
    def _seg(arr, f, t):
        s = arr[f:t]
        return np.array(s, copy=True) if isinstance(s, np.ndarray) else list(s)

    def _generic_array(self, arr, operand):
        """Normalises host arrays for the generic path: numpy view or the python list itself."""
        if _is_torch(arr):
            return self._host_view(arr, operand)
        return arr

    # ------------------------------------------------------------------ gather
    def gatherArray(self, arrData, operand, sendfroms, sendtos, rootRank):
        """Rank i's [from_i, to_i) lands at the same offsets in root's array (PCS:535)."""
        CommUtils.isfromsTosLegal(sendfroms, sendtos, size=len(arrData))
        self._check_root(rootRank)
        if self._n == 1:
            return arrData
        if _is_cuda(arrData):
            plane = self._cuda_plane()
            if plane is None:
                return self._staged(arrData, lambda h: self.gatherArray(h, operand, sendfroms, sendtos, rootRank))
            plane.gather(arrData, operand, sendfroms, sendtos, rootRank)
            return arrData
        a = self._native_host(arrData, operand, None)
        if a is not None:
            check(N.lib.mp4j_host_gather(self._h, a.ctypes.data, operand.code, N.i64_array(sendfroms),
                                         N.i64_array(sendtos), rootRank))
            return arrData
        g = self._generic_array(arrData, operand)
        r = self._rank
        acc = self._tree_gather(self._seg(g, sendfroms[r], sendtos[r]), rootRank, operand)
        if r == rootRank:
            for q, seg in acc.items():
                if q != r:
                    g[sendfroms[q]:sendtos[q]] = seg
        return arrData

    def gatherMap(self, mapData, operand, rootRank):
        """Root receives the union of all maps (duplicate keys: last writer wins, PCS:477-481)."""
        self._check_root(rootRank)
        if self._n == 1:
            return mapData
        acc = self._tree_gather(mapData, rootRank, self._map_wire(operand))
        if self._rank != rootRank:
            return {}
        out = {}
        for q in sorted(acc):
            out.update(acc[q])
        return out

    # ------------------------------------------------------------------ allgather
    def allgatherArray(self, arrData, operand, froms, tos):
        CommUtils.isfromsTosLegal(froms, tos, size=len(arrData))
        if self._n == 1:
            return arrData
        if _is_cuda(arrData):
            plane = self._cuda_plane()
            if plane is None:
                return self._staged(arrData, lambda h: self.allgatherArray(h, operand, froms, tos))
            plane.allgather(arrData, operand, froms, tos)
            return arrData
        a = self._native_host(arrData, operand, None)
        if a is not None:
            check(N.lib.mp4j_host_allgather(self._h, a.ctypes.data, operand.code, N.i64_array(froms),
                                            N.i64_array(tos)))
            return arrData
        g = self._generic_array(arrData, operand)
        r = self._rank
        parts = self._ring_allgather(self._seg(g, froms[r], tos[r]), operand)
        for q in range(self._n):
            if q != r:
                g[froms[q]:tos[q]] = parts[q]
        return arrData

    def allgatherMap(self, mapData, operand):
        """Returns the list of p maps indexed by origin rank (PCS:666, 714-715)."""
        if self._n == 1:
            return [mapData]
        if self._gpu_map_ok(operand, None, need_op=False):
            out = self._gpu_allgather_map([mapData], operand)
            if out is not None:
                return out
        return self._ring_allgather(mapData, self._map_wire(operand))

    # ------------------------------------------------------------------ broadcast
    def broadcastArray(self, arrData, operand, from_, to, rootRank):
        """= scatter of p balanced chunks + allgather (PCS:750-770)."""
        CommUtils.isFromToLegal(from_, to, len(arrData))
        self._check_root(rootRank)
        if self._n == 1:
            return arrData
        if _is_cuda(arrData):
            plane = self._cuda_plane()
            if plane is None:
                return self._staged(arrData, lambda h: self.broadcastArray(h, operand, from_, to, rootRank))
            plane.broadcast(arrData, operand, from_, to, rootRank)
            return arrData
        a = self._native_host(arrData, operand, None)
        if a is not None:
            check(N.lib.mp4j_host_broadcast(self._h, a.ctypes.data, operand.code, from_, to, rootRank))
            return arrData
        froms, tos = CommUtils.splitRange(from_, to, self._n)
        self.scatterArray(arrData, operand, froms, tos, rootRank)
        return self.allgatherArray(arrData, operand, froms, tos)

    def broadcast(self, value, operand, rootRank):
        """Scalar broadcast: boxed into a 1-element array (PCS:786-829)."""
        if self._n == 1:
            return value
        box = self._box(value, operand)
        self.broadcastArray(box, operand, 0, 1, rootRank)
        return self._unbox(box)

    def broadcastMap(self, mapData, operand, rootRank):
        """Root's map reaches everyone: hash-partition + scatter + allgather + merge (PCS:842-880)."""
        self._check_root(rootRank)
        if self._n == 1:
            return mapData
        parts = None
        if self._rank == rootRank:
            parts = [dict() for _ in range(self._n)]
            for k, v in mapData.items():
                parts[key_partition(k, self._n)][k] = v
        wire = self._map_wire(operand)
        mine = self._tree_scatter(parts, rootRank, wire)
        out = {}
        for m in self._ring_allgather(mine, wire):
            out.update(m)
        return out

    # ------------------------------------------------------------------ scatter
    def scatterArray(self, arrData, operand, recvfroms, recvtos, rootRank):
        CommUtils.isfromsTosLegal(recvfroms, recvtos, size=len(arrData))
        self._check_root(rootRank)
        if self._n == 1:
            return arrData
        if _is_cuda(arrData):
            plane = self._cuda_plane()
            if plane is None:
                return self._staged(arrData, lambda h: self.scatterArray(h, operand, recvfroms, recvtos, rootRank))
            plane.scatter(arrData, operand, recvfroms, recvtos, rootRank)
            return arrData
        a = self._native_host(arrData, operand, None)
        if a is not None:
            check(N.lib.mp4j_host_scatter(self._h, a.ctypes.data, operand.code, N.i64_array(recvfroms),
                                          N.i64_array(recvtos), rootRank))
            return arrData
        g = self._generic_array(arrData, operand)
        r = self._rank
        payloads = None
        if r == rootRank:
            payloads = [self._seg(g, recvfroms[q], recvtos[q]) for q in range(self._n)]
        mine = self._tree_scatter(payloads, rootRank, operand)
        if r != rootRank:
            g[recvfroms[r]:recvtos[r]] = mine
        return arrData

    def scatterMap(self, mapDataList, operand, rootRank):
        """Root holds p maps; rank i receives map i (PCS:949)."""
        self._check_root(rootRank)
        if self._n == 1:
            return mapDataList[0]
        if self._rank == rootRank and len(mapDataList) != self._n:
            raise Mp4jException("scatterMap needs one map per slave on the root")
        return self._tree_scatter(mapDataList if self._rank == rootRank else None, rootRank, self._map_wire(operand))

    def _scatterMapSpecial(self, mapListList, operand, rootRank):
        """T maps per rank (used by ThreadCommSlave.scatterMap, PCS:999)."""
        if self._n == 1:
            return mapListList[0]
        return self._tree_scatter(mapListList if self._rank == rootRank else None, rootRank, operand)

    # ------------------------------------------------------------------ reduce-scatter
    def reduceScatterArray(self, arrData, operand, operator, from_, counts):
        """Rank r ends with the fully reduced [from+sum(counts[:r]), +counts[r]) (PCS:1174).
        Outside the owned slice the array keeps its input values on the GPU/generic paths
        (the reference leaves partial sums there, PCS:1162-1172; both are 'unspecified')."""
        CommUtils.isFromCountsLegal(from_, counts, len(arrData))
        if len(counts) != self._n:
            raise Mp4jException("counts must have one entry per slave")
        if self._n == 1:
            return arrData
        if _is_cuda(arrData):
            plane = self._cuda_plane()
            if plane is None:
                return self._staged(arrData, lambda h: self.reduceScatterArray(h, operand, operator, from_, counts),
                                    operator)
            plane.reduce_scatter(arrData, operand, operator, from_, counts)
            return arrData
        a = self._native_host(arrData, operand, operator)
        if a is not None:
            check(N.lib.mp4j_host_reduce_scatter(self._h, a.ctypes.data, operand.code, operator.code, from_,
                                                 N.i64_array(counts)))
            return arrData
        g = self._generic_array(arrData, operand)
        froms = CommUtils.getFromsFromCount(from_, counts, self._n)
        tos = CommUtils.getTosFromCount(from_, counts, self._n)
        parts = [self._seg(g, froms[q], tos[q]) for q in range(self._n)]
        mine = self._ring_reduce_scatter(parts, self._merge_seq(operator, operand), operand)
        g[froms[self._rank]:tos[self._rank]] = mine
        return arrData

    def reduceScatterMap(self, mapDataList, operand, operator):
        """mapDataList[i] = this rank's contribution to rank i; returns the reduced map i (PCS:1299)."""
        if self._n == 1:
            return mapDataList[0]
        if len(mapDataList) != self._n:
            raise Mp4jException("reduceScatterMap needs one map per slave")
        if self._gpu_map_ok(operand, operator):
            out = self._gpu_reduce_scatter_map(list(enumerate(mapDataList)), operand, operator)
            if out is not None:
                return out
        if self._native_map_ok(operand, operator) and all(isinstance(m, dict) for m in mapDataList):
            return self._native_map_reduce_scatter(mapDataList, operand, operator)
        return self._ring_reduce_scatter(list(mapDataList), self._merge_map(operator), operand)

    # ------------------------------------------------------------------ reduce
    def reduceArray(self, arrData, operand, operator, from_, to, rootRank):
        """= reduce-scatter + gather (PCS:1390-1416)."""
        CommUtils.isFromToLegal(from_, to, len(arrData))
        self._check_root(rootRank)
        if self._n == 1:
            return arrData
        if _is_cuda(arrData):
            plane = self._cuda_plane()
            if plane is None:
                return self._staged(arrData, lambda h: self.reduceArray(h, operand, operator, from_, to, rootRank),
                                    operator)
            plane.reduce(arrData, operand, operator, from_, to, rootRank)
            return arrData
        a = self._native_host(arrData, operand, operator)
        if a is not None:
            check(N.lib.mp4j_host_reduce(self._h, a.ctypes.data, operand.code, operator.code, from_, to, rootRank))
            return arrData
        froms, tos = CommUtils.splitRange(from_, to, self._n)
        self.reduceScatterArray(arrData, operand, operator, from_, [t - f for f, t in zip(froms, tos)])
        return self.gatherArray(arrData, operand, froms, tos, rootRank)

    def reduce(self, value, operand, operator, rootRank):
        if self._n == 1:
            return value
        box = self._box(value, operand)
        self.reduceArray(box, operand, operator, 0, 1, rootRank)
        return self._unbox(box)

    def reduceMap(self, mapData, operand, operator, rootRank):
        """hash-partition + reduce-scatter + gather to root (PCS:1490-1516)."""
        self._check_root(rootRank)
        if self._n == 1:
            return mapData
        if self._native_map_ok(operand, operator):
            return self._native_map_reduce(1, rootRank, mapData, operand, operator)
        mine = self.reduceScatterMap(self._partition_map(mapData), operand, operator)
        return self.gatherMap(mine, operand, rootRank)

    # ------------------------------------------------------------------ allreduce
    def allreduceArray(self, arrData, operand, operator, from_=0, to=None):
        """= reduce-scatter + allgather on the host paths (PCS:1733-1763); on the GPU one fused
        kernel chosen by size (one-shot / two-shot / NVLS)."""
        if to is None:
            to = len(arrData)
        CommUtils.isFromToLegal(from_, to, len(arrData))
        if self._n == 1:
            return arrData
        if _is_cuda(arrData):
            plane = self._cuda_plane()
            if plane is None:
                return self._staged(arrData, lambda h: self.allreduceArray(h, operand, operator, from_, to), operator)
            plane.allreduce(arrData, operand, operator, from_, to)
            return arrData
        if self._gpu_host_ok(arrData, operand, operator):
            self._gpu.thread_allreduce_host([arrData], operand, operator, from_, to)
            return arrData
        a = self._native_host(arrData, operand, operator)
        if a is not None:
            check(N.lib.mp4j_host_allreduce(self._h, a.ctypes.data, operand.code, operator.code, from_, to))
            return arrData
        froms, tos = CommUtils.splitRange(from_, to, self._n)
        self.reduceScatterArray(arrData, operand, operator, from_, [t - f for f, t in zip(froms, tos)])
        return self.allgatherArray(arrData, operand, froms, tos)

    def allreduce(self, value, operand, operator):
        if self._n == 1:
            return value
        box = self._box(value, operand)
        self.allreduceArray(box, operand, operator, 0, 1)
        return self._unbox(box)

    def allreduceMap(self, mapData, operand, operator):
        """Keys are partitioned by String.hashCode % slaveNum, reduce-scattered, allgathered and
        merged (PCS:2053-2081). For tensor-backed sparse maps use gpu.allreduce_map."""
        if self._n == 1:
            return mapData
        if self._gpu_map_ok(operand, operator):
            out = self._gpu_allreduce_map([mapData], operand, operator)
            if out is not None:
                return out
        if self._native_map_ok(operand, operator):
            return self._native_map_reduce(0, 0, mapData, operand, operator)
        mine = self.reduceScatterMap(self._partition_map(mapData), operand, operator)
        out = {}
        for m in self.allgatherMap(mine, operand):
            out.update(m)
        return out

    # ------------------------------------------------------------------ RPC allreduce (small data)
    def allreduceArrayRpc(self, arrData, operand, operator):
        """Every slave ships the whole array to the master mailbox and reduces the p copies
        locally (PCS:1776-1799, SRV:373-497). On CUDA tensors this is the one-shot kernel, which
        has the same role: small arrays where set-up cost dominates."""
        if self._n == 1:
            return arrData
        if _is_cuda(arrData):
            plane = self._cuda_plane()
            if plane is None:
                return self._staged(arrData, lambda h: self.allreduceArrayRpc(h, operand, operator), operator)
            plane.allreduce(arrData, operand, operator, 0, arrData.numel(), algo=N.ALGO_ONESHOT)
            return arrData
        a = self._native_host(arrData, operand, operator)
        if a is not None:
            check(N.lib.mp4j_host_allreduce_rpc(self._h, a.ctypes.data, operand.code, operator.code, a.size))
            return arrData
        g = self._generic_array(arrData, operand)
        data = operand.dumps(self._seg(g, 0, len(g)))
        ptr, n = C.c_void_p(), C.c_size_t()
        check(N.lib.mp4j_comm_gather_all(self._h, data, len(data), C.byref(ptr), C.byref(n)))
        raw = N.take_bytes(ptr, n)
        merge = self._merge_seq(operator, operand)
        acc, off = None, 0
        for _ in range(self._n):
            ln = int.from_bytes(raw[off:off + 8], "little")
            part = operand.loads(raw[off + 8:off + 8 + ln])
            off += 8 + ln
            acc = part if acc is None else merge(acc, part)
        g[0:len(g)] = acc
        return arrData

    def allreduceRpc(self, value, operand, operator):
        if self._n == 1:
            return value
        box = self._box(value, operand)
        self.allreduceArrayRpc(box, operand, operator)
        return self._unbox(box)

    # ------------------------------------------------------------------ set / list helpers (PCS:1583-1718, 2099-2211)
    @staticmethod
    def _obj_operand(elementSerializer=None, elementType=None):
        return Operands.OBJECT_OPERAND(elementSerializer, elementType)

    _UNION = custom(lambda a, b: a | b, "SET_UNION")
    _INTERSECT = custom(lambda a, b: a & b, "SET_INTERSECTION")
    _CONCAT = custom(lambda a, b: a + b, "LIST_CONCAT")

    def reduceMapSetUnion(self, mapData, rootRank, elementSerializer=None, elementType=None):
        return self.reduceMap(mapData, self._obj_operand(elementSerializer, elementType), self._UNION, rootRank)

    def reduceSetUnion(self, setData, rootRank, elementSerializer=None, elementType=None):
        return self.reduceMapSetUnion({"key": setData}, rootRank, elementSerializer, elementType).get("key")

    def reduceMapSetIntersection(self, mapData, rootRank, elementSerializer=None, elementType=None):
        return self.reduceMap(mapData, self._obj_operand(elementSerializer, elementType), self._INTERSECT, rootRank)

    def reduceSetIntersection(self, setData, rootRank, elementSerializer=None, elementType=None):
        return self.reduceMapSetIntersection({"key": setData}, rootRank, elementSerializer, elementType).get("key")

    def reduceMapListConcat(self, mapData, rootRank, elementSerializer=None, elementType=None):
        return self.reduceMap(mapData, self._obj_operand(elementSerializer, elementType), self._CONCAT, rootRank)

    def reduceListConcat(self, listData, rootRank, elementSerializer=None, elementType=None):
        return self.reduceMapListConcat({"key": listData}, rootRank, elementSerializer, elementType).get("key")

    def allreduceMapSetUnion(self, mapData, elementSerializer=None, elementType=None):
        return self.allreduceMap(mapData, self._obj_operand(elementSerializer, elementType), self._UNION)

    def allreduceSetUnion(self, setData, elementSerializer=None, elementType=None):
        return self.allreduceMapSetUnion({"key": setData}, elementSerializer, elementType)["key"]

    def allreduceMapSetIntersection(self, mapData, elementSerializer=None, elementType=None):
        return self.allreduceMap(mapData, self._obj_operand(elementSerializer, elementType), self._INTERSECT)

    def allreduceSetIntersection(self, setData, elementSerializer=None, elementType=None):
        return self.allreduceMapSetIntersection({"key": setData}, elementSerializer, elementType)["key"]

    def allreduceMapListConcat(self, mapData, elementSerializer=None, elementType=None):
        return self.allreduceMap(mapData, self._obj_operand(elementSerializer, elementType), self._CONCAT)

    def allreduceListConcat(self, listData, elementSerializer=None, elementType=None):
        return self.allreduceMapListConcat({"key": listData}, elementSerializer, elementType)["key"]

    # ------------------------------------------------------------------ native map path
    # ------------------------------------------------------------------ Map<String, number> on the GPU
    def _map_bridge(self):
        if getattr(self, "_bridge", None) is None:
            from .map_bridge import MapKeyBridge
            self._bridge = MapKeyBridge(self)
        return self._bridge

    def _gpu_map_ok(self, operand, operator, need_op=True):
        """Maps with float/double/int/long values and SUM/MAX/MIN go through the GPU map kernels once the GPU
        plane is up (String keys -> 64-bit keys: comm/map_bridge.py). Depends only on state that is the same on
        every rank; MP4J_GPU_MAPS=0 keeps maps on the host."""
        if self._gpu is None or self._n == 1 or os.environ.get("MP4J_GPU_MAPS", "1") == "0":
            return False
        if not operand.is_primitive or operand.code not in (N.F64, N.F32, N.I64, N.I32):
            return False
        if need_op and (operator is None or not operator.is_native or operator.code not in (N.SUM, N.MAX, N.MIN)):
            return False
        return not self._map_bridge().broken

    def _gpu_map_tensors(self, maps, operand, tags=None):
        """(keys, values) CUDA tensors of the concatenation of `maps` (a list of dicts); None if the bridge broke.
        tags: destination rank per map (partition "tagged")."""
        import torch
        br = self._map_bridge()
        strings = [k for m in maps for k in m.keys()]
        ks = br.encode(strings)
        if br.broken:
            return None
        if tags is not None:
            off = 0
            for m, d in zip(maps, tags):
                ks[off:off + len(m)] = br.tag(ks[off:off + len(m)], d)
                off += len(m)
        vals = np.fromiter((v for m in maps for v in m.values()), dtype=operand.np_dtype, count=len(strings))
        dev = "cuda:%d" % self._gpu.device
        return torch.from_numpy(ks).to(dev), torch.from_numpy(vals).to(dev)

    def _gpu_map_result(self, k, v, operand, tagged=False):
        import torch
        torch.cuda.current_stream().synchronize()
        names = self._map_bridge().decode(k.cpu().numpy().tolist(), tagged)
        return dict(zip(names, v.cpu().numpy().tolist()))

    def _gpu_allreduce_map(self, maps, operand, operator):
        kv = self._gpu_map_tensors(maps, operand)
        if kv is None:
            return None
        k, v = self._gpu.allreduce_map(kv[0], kv[1], operator, partition="java32")
        return self._gpu_map_result(k, v, operand)

    def _gpu_reduce_scatter_map(self, maps_by_dest, operand, operator):
        """maps_by_dest: list of (dest rank, dict); returns this rank's reduced dict."""
        kv = self._gpu_map_tensors([m for _, m in maps_by_dest], operand, tags=[d for d, _ in maps_by_dest])
        if kv is None:
            return None
        k, v = self._gpu.reduce_scatter_map(kv[0], kv[1], operator, partition="tagged")
        return self._gpu_map_result(k, v, operand, tagged=True)

    def _gpu_allgather_map(self, maps, operand):
        """returns the list of every rank's (merged) dict, in rank order."""
        kv = self._gpu_map_tensors(maps, operand)
        if kv is None:
            return None
        lop = Operands.LONG_OPERAND()
        counts = np.zeros(self._n, dtype=np.int64)
        counts[self._rank] = kv[0].numel()
        self.allgatherArray(counts, lop, list(range(self._n)), list(range(1, self._n + 1)))
        k, v = self._gpu.allgather_map(kv[0], kv[1])
        import torch
        torch.cuda.current_stream().synchronize()
        names = self._map_bridge().decode(k.cpu().numpy().tolist())
        vals = v.cpu().numpy().tolist()
        out, off = [], 0
        for c in counts.tolist():
            out.append(dict(zip(names[off:off + c], vals[off:off + c])))
            off += c
        return out

    def _map_wire(self, operand):
        """Codec of map payloads on the move-only map collectives: the packed native format for primitive values."""
        if operand.is_primitive and operand.np_dtype is not None and os.environ.get("MP4J_NATIVE_MAPS", "1") != "0":
            return _MapCodec(operand)
        return operand

    def _native_map_reduce_scatter(self, mapDataList, operand, operator):
        """reduceScatterMap natively (csrc/host/map_host.cc): the p maps back to back, ring reduce-scatter with the
        native string-keyed merge."""
        counts = N.i64_array([len(m) for m in mapDataList])
        n = sum(len(m) for m in mapDataList)
        try:
            blob = "\0".join(k for m in mapDataList for k in m.keys()).encode("utf-8")
        except TypeError:
            raise Mp4jException("map keys must be str (Map<String, V> in the reference)")
        if n and blob.count(b"\0") != n - 1:
            raise Mp4jException("map keys must not contain NUL characters")
        vals = np.fromiter((v for m in mapDataList for v in m.values()), dtype=operand.np_dtype, count=n)
        ok, ov = C.c_void_p(), C.c_void_p()
        okl, on = C.c_size_t(), C.c_size_t()
        check(N.lib.mp4j_host_map_reduce_scatter(self._h, blob, len(blob), vals.ctypes.data, counts, operand.code,
                                                 operator.code, C.byref(ok), C.byref(okl), C.byref(ov), C.byref(on)))
        m = on.value
        keys_raw = N.take_bytes(ok, okl)
        vals_raw = N.take_bytes(ov, C.c_size_t(m * np.dtype(operand.np_dtype).itemsize))
        if m == 0:
            return {}
        return dict(zip(keys_raw.decode("utf-8").split("\0"),
                        np.frombuffer(vals_raw, dtype=operand.np_dtype, count=m).tolist()))

    @staticmethod
    def _native_map_ok(operand, operator):
        """Maps with primitive values and a predefined operator take the C++ map collective
        (csrc/host/map_host.cc). The choice depends on (operand, operator) only, so that every rank
        takes the same path; MP4J_NATIVE_MAPS=0 forces the generic Python ring everywhere."""
        return (operand.is_primitive and operand.np_dtype is not None and operator is not None
                and operator.is_native and os.environ.get("MP4J_NATIVE_MAPS", "1") != "0")

    def _native_map_reduce(self, mode, root, mapData, operand, operator):
        n = len(mapData)
        try:
            blob = "\0".join(mapData.keys()).encode("utf-8")
        except TypeError:
            raise Mp4jException("map keys must be str (Map<String, V> in the reference)")
        if n and blob.count(b"\0") != n - 1:
            raise Mp4jException("map keys must not contain NUL characters")
        vals = np.fromiter(mapData.values(), dtype=operand.np_dtype, count=n)
        ok, ov = C.c_void_p(), C.c_void_p()
        okl, on = C.c_size_t(), C.c_size_t()
        check(N.lib.mp4j_host_map_reduce(self._h, mode, root, blob, len(blob), vals.ctypes.data, n, operand.code,
                                         operator.code, C.byref(ok), C.byref(okl), C.byref(ov), C.byref(on)))
        m = on.value
        keys_raw = N.take_bytes(ok, okl)
        vals_raw = N.take_bytes(ov, C.c_size_t(m * np.dtype(operand.np_dtype).itemsize))
        if m == 0:
            return {}
        out_vals = np.frombuffer(vals_raw, dtype=operand.np_dtype, count=m).tolist()
        return dict(zip(keys_raw.decode("utf-8").split("\0"), out_vals))

    # ------------------------------------------------------------------ internals
    def _check_root(self, root):
        if root < 0 or root >= self._n:
            raise Mp4jException("rootRank %d out of range [0,%d)" % (root, self._n))

    def _partition_map(self, mapData):
        parts = [dict() for _ in range(self._n)]
        p = self._n
        for k, v in mapData.items():
            parts[key_partition(k, p)][k] = v
        return parts

    @staticmethod
    def _box(value, operand):
        if operand.is_primitive and operand.np_dtype is not None:
            return np.array([value]).astype(operand.np_dtype, casting="unsafe")
        return [value]

    @staticmethod
    def _unbox(box):
        v = box[0]
        return v.item() if isinstance(v, np.generic) else v

    # snake_case aliases
    get_rank = getRank
    get_slave_num = getSlaveNum
    gather_array, gather_map = gatherArray, gatherMap
    allgather_array, allgather_map = allgatherArray, allgatherMap
    broadcast_array, broadcast_map = broadcastArray, broadcastMap
    scatter_array, scatter_map = scatterArray, scatterMap
    reduce_scatter_array, reduce_scatter_map = reduceScatterArray, reduceScatterMap
    reduce_array, reduce_map = reduceArray, reduceMap
    allreduce_array, allreduce_map = allreduceArray, allreduceMap
    allreduce_array_rpc, allreduce_rpc = allreduceArrayRpc, allreduceRpc
    write_file = writeFile
