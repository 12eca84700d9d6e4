"""Python face of the GPU data plane (csrc/gpu, include/mp4j_gpu.h).

GpuPlane  - attached to a ProcessCommSlave: one process per GPU, peers mapped over NVLink.
EmuPlane  - `world` emulated ranks on ONE GPU (every collective is one cooperative launch whose
            grid rows are the ranks): how the kernels' multi-rank logic is tested on a single GPU.
Both only marshal pointers: all work happens in libmp4j.so. There is no eager/PyTorch fallback;
a missing or failing native call raises Mp4jException.
"""
import ctypes as C
import os

import numpy as np
import torch

from .. import _native as N
from ..exception import Mp4jException, check

_P = N._proto
_P("mp4j_gpu_init", C.c_int, C.c_void_p, C.c_int, C.c_size_t, C.c_size_t, N.c_vpp)
_P("mp4j_gpu_emu_create", C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, N.c_vpp)
_P("mp4j_gpu_destroy", None, C.c_void_p)
_P("mp4j_gpu_world", C.c_int, C.c_void_p)
_P("mp4j_gpu_rank", C.c_int, C.c_void_p)
_P("mp4j_gpu_has_multicast", C.c_int, C.c_void_p)
_P("mp4j_gpu_describe", C.c_int, C.c_void_p, C.c_char_p, C.c_size_t)
_P("mp4j_gpu_last", C.c_int, C.c_void_p, C.c_char_p, C.c_size_t)
_P("mp4j_gpu_alloc", C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t))
_P("mp4j_gpu_free", C.c_int, C.c_void_p, C.c_size_t)
_P("mp4j_gpu_ptr", C.c_void_p, C.c_void_p, C.c_int, C.c_size_t)
_P("mp4j_gpu_check", C.c_int, C.c_void_p)
_P("mp4j_gpu_allreduce", C.c_int, C.c_void_p, N.c_vpp, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_void_p)
_P("mp4j_gpu_reduce_scatter", C.c_int, C.c_void_p, N.c_vpp, C.c_int, C.c_int, C.c_int64, N.c_i64p, C.c_int,
   C.c_void_p)
_P("mp4j_gpu_allgather", C.c_int, C.c_void_p, N.c_vpp, C.c_int, N.c_i64p, N.c_i64p, C.c_int, C.c_void_p)
_P("mp4j_gpu_broadcast", C.c_int, C.c_void_p, N.c_vpp, C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_void_p)
_P("mp4j_gpu_gather", C.c_int, C.c_void_p, N.c_vpp, C.c_int, N.c_i64p, N.c_i64p, C.c_int, C.c_void_p)
_P("mp4j_gpu_scatter", C.c_int, C.c_void_p, N.c_vpp, C.c_int, N.c_i64p, N.c_i64p, C.c_int, C.c_void_p)
_P("mp4j_gpu_reduce", C.c_int, C.c_void_p, N.c_vpp, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_int,
   C.c_void_p)
_P("mp4j_gpu_thread_allreduce", C.c_int, C.c_void_p, N.c_vpp, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64,
   C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p)

_P("mp4j_gpu_thread_segments", C.c_int, C.c_void_p, N.c_vpp, C.c_int, C.c_int, C.c_int, N.c_i64p, C.c_int, C.c_char_p,
   C.c_void_p)
_P("mp4j_gpu_thread_allreduce_host", C.c_int, C.c_void_p, N.c_vpp, C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int64,
   C.c_int, C.c_int, C.c_void_p)

_P("mp4j_gpu_map_collective", C.c_int, C.c_void_p, C.c_int, N.c_vpp, N.c_vpp, N.c_i64p, C.c_int64, C.c_int64, C.c_int,
   C.c_int, C.c_int, N.c_vpp, N.c_vpp, N.c_vpp, C.c_void_p)

_P("mp4j_gpu_map_reduce_local", C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int,
   C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p)

_P("mp4j_gpu_lr_xtr_reduce_scatter", C.c_int, C.c_void_p, N.c_vpp, N.c_vpp, C.c_int, C.c_int, C.c_int64, N.c_vpp,
   C.POINTER(C.c_int), C.c_void_p)

_P("mp4j_gpu_gbdt_hist_reduce_scatter", C.c_int, C.c_void_p, N.c_vpp, N.c_i64p, N.c_vpp, N.c_vpp, N.c_i64p, C.c_int,
   N.c_vpp, C.POINTER(C.c_int), C.c_void_p)

_P("mp4j_gpu_lr_softmax_residual", C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p)
_P("mp4j_gpu_allreduce_wire", C.c_int, C.c_void_p, N.c_vpp, C.c_int64, C.c_int64, C.c_int, C.c_void_p)
_P("mp4j_gpu_custom_op", C.c_int, C.c_void_p, C.c_int, C.c_char_p, C.POINTER(C.c_int))
_P("mp4j_gpu_trace", C.c_int, C.c_void_p, C.c_int)
_P("mp4j_gpu_set_option", C.c_int, C.c_void_p, C.c_char_p, C.c_longlong)
_P("mp4j_thread_gpu_allreduce", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int64,
   C.c_int64, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p)
_P("mp4j_thread_gpu_allreduce_host", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int64,
   C.c_int64, C.c_int, C.c_int, C.c_void_p)
_P("mp4j_thread_gpu_segments", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
   C.POINTER(C.c_int64), C.c_int, C.c_char_p, C.c_void_p)
_P("mp4j_gpu_stats", C.c_int, C.c_void_p, C.c_char_p, C.c_size_t, C.c_int)
_P("mp4j_rtc_compile", C.c_int, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_longlong))

MAP_MODES = {"allreduce": 0, "reduce_scatter": 1, "allgather": 2}
MAP_PARTITIONS = {"mix": 0, "java32": 1, "tagged": 2}

TORCH_DTYPES = {
    N.F64: torch.float64, N.F32: torch.float32, N.I64: torch.int64, N.I32: torch.int32,
    N.I16: torch.int16, N.I8: torch.int8, N.BF16: torch.bfloat16, N.F16: torch.float16,
}
DTYPE_CODES = {v: k for k, v in TORCH_DTYPES.items()}
_TYPESTR = {N.F64: "<f8", N.F32: "<f4", N.I64: "<i8", N.I32: "<i4", N.I16: "<i2", N.I8: "|i1", N.F16: "<f2"}
_ES = {N.F64: 8, N.F32: 4, N.I64: 8, N.I32: 4, N.I16: 2, N.I8: 1, N.BF16: 2, N.F16: 2}

# intra-GPU thread-reduce variants: auto = 256-bit static-T fast path when eligible, else register
THREAD_ALGO = {"auto": 0, "cluster": 1, "register": 3}


def _apply_tuning_table(world):
    """K0, runtime half: AUTO's thresholds come from the measured table tuning/b200_p<world>.json
    (written by tools/make_tuning.py from a bench/sweep.py run). Explicit environment variables win."""
    import json
    path = os.environ.get("MP4J_TUNE_FILE") or os.path.join(
        os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "tuning", "b200_p%d.json" % world)
    try:
        table = json.load(open(path))
    except Exception:
        return None
    for key, env in (("oneshot_max_bytes", "MP4J_ONE_SHOT_MAX_BYTES"), ("nvls_min_bytes", "MP4J_NVLS_MIN_BYTES"),
                     ("nvls_min_world", "MP4J_NVLS_MIN_WORLD"), ("rs_nvls_min_bytes", "MP4J_RS_NVLS_MIN_BYTES"),
                     ("ag_nvls_min_bytes", "MP4J_AG_NVLS_MIN_BYTES"), ("rs_nvls_max_bytes", "MP4J_RS_NVLS_MAX_BYTES"),
                     ("ag_nvls_max_bytes", "MP4J_AG_NVLS_MAX_BYTES")):
        if key in table and env not in os.environ:
            os.environ[env] = str(int(min(table[key], 1 << 62)))
    return table


class _RawCuda:
    """Minimal __cuda_array_interface__ carrier so torch can wrap symmetric-heap memory."""

    def __init__(self, ptr, n, code, owner):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": _TYPESTR[code], "data": (ptr, False),
                                         "version": 2, "strides": None}
        self._owner = owner


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def _ptr_array(tensors):
    return (C.c_void_p * len(tensors))(*[C.c_void_p(t.data_ptr()) for t in tensors])


def _code_of(t, operand=None):
    code = DTYPE_CODES.get(t.dtype)
    if code is None:
        raise Mp4jException("unsupported tensor dtype %s" % t.dtype)
    if operand is not None and operand.code != code:
        raise Mp4jException("tensor dtype %s does not match operand %s" % (t.dtype, operand.name))
    return code


def _check_tensor(t):
    if not t.is_cuda or t.dim() != 1 or not t.is_contiguous():
        raise Mp4jException("GPU collectives need 1-D contiguous CUDA tensors")


def _check_host_tensor(t):
    if t.is_cuda or t.dim() != 1 or not t.is_contiguous():
        raise Mp4jException("the host-array pipeline needs 1-D contiguous CPU tensors (pinned for full speed)")


def lr_softmax_residual(logits, y, rt=None, loss=None, stream=None):
    """Fused softmax residual for LR (csrc/gpu/lr_residual.cu): logits bf16 [rows, C], y int64 [rows] (< 0 = padding).
    Returns (Rt bf16 [C, rows] = softmax - onehot, loss fp32 [1] incremented by the summed NLL)."""
    rows, Cn = logits.shape
    if logits.dtype != torch.bfloat16 or not logits.is_contiguous() or y.dtype != torch.int64:
        raise Mp4jException("lr_softmax_residual needs contiguous bf16 logits and int64 labels")
    if rt is None:
        rt = torch.empty(Cn, rows, dtype=torch.bfloat16, device=logits.device)
    if loss is None:
        loss = torch.zeros(1, dtype=torch.float32, device=logits.device)
    check(N.lib.mp4j_gpu_lr_softmax_residual(logits.data_ptr(), y.data_ptr(), rows, Cn, rt.data_ptr(), loss.data_ptr(),
                                             _stream_ptr(stream)))
    return rt, loss


def _native_op(operator):
    if operator is None:
        return N.COPY
    if not operator.is_native:
        raise Mp4jException("custom Python operators run on the host path only; move the tensor to the CPU, "
                            "use a predefined operator, or (dense array collectives) Operators.device_expr")
    return operator.code


class _PlaneBase:
    """Shared marshalling; subclasses define self._g and how many local ranks an `arrs` list has."""

    _g = None

    def describe(self):
        buf = C.create_string_buffer(1024)
        check(N.lib.mp4j_gpu_describe(self._g, buf, 1024))
        return buf.value.decode()

    def last(self):
        buf = C.create_string_buffer(512)
        check(N.lib.mp4j_gpu_last(self._g, buf, 512))
        return buf.value.decode()

    @property
    def world(self):
        return N.lib.mp4j_gpu_world(self._g)

    @property
    def has_multicast(self):
        return bool(N.lib.mp4j_gpu_has_multicast(self._g))

    def _op(self, operator, code):
        """Native op code of `operator` for element type `code` on this plane; device expressions
        (Operators.device_expr) are compiled and loaded on first use."""
        if operator is not None and operator.code is None and getattr(operator, "expr", None):
            return self.custom_op(operator.expr, code)
        return _native_op(operator)

    def custom_op(self, expr, code):
        """Compiles (NVRTC, sm_100a) and loads the device expression `expr` for element type `code`;
        returns the op code the dense reducing collectives of this plane accept."""
        cache = self.__dict__.setdefault("_custom_ops", {})
        key = (code, expr)
        if key not in cache:
            out = C.c_int()
            check(N.lib.mp4j_gpu_custom_op(self._g, code, expr.encode(), C.byref(out)))
            cache[key] = out.value
        return cache[key]

    def trace(self, on=True):
        """Per-collective CUDA-event timing + NVTX ranges on/off (also MP4J_TRACE=1)."""
        check(N.lib.mp4j_gpu_trace(self._g, 1 if on else 0))

    def stats(self, reset=False):
        """Per (collective, algorithm) statistics since tracing was enabled: calls, bytes, total/min/
        max/avg microseconds on the device, algorithmic GB/s. Synchronises the traced events."""
        import json
        buf = C.create_string_buffer(1 << 16)
        check(N.lib.mp4j_gpu_stats(self._g, buf, len(buf), 1 if reset else 0))
        return json.loads(buf.value.decode())

    def set_option(self, name, value):
        """Run-time tuning knob (sweeps). Every rank must set the same value before the next collective."""
        check(N.lib.mp4j_gpu_set_option(self._g, name.encode(), int(value)))

    def check(self):
        """Raises if a kernel hit the device-side peer-wait watchdog since the last call."""
        check(N.lib.mp4j_gpu_check(self._g))

    def _alloc_off(self, nbytes):
        off = C.c_size_t()
        check(N.lib.mp4j_gpu_alloc(self._g, nbytes, C.byref(off)))
        return off.value

    def _wrap(self, rank, off, n, code):
        ptr = N.lib.mp4j_gpu_ptr(self._g, rank, off)
        if code == N.BF16:
            t = torch.as_tensor(_RawCuda(ptr, n, N.I16, self), device="cuda:%d" % self.device)
            return t.view(torch.bfloat16)
        return torch.as_tensor(_RawCuda(ptr, n, code, self), device="cuda:%d" % self.device)

    def ws_trace(self, rank=None):
        """Diagnostics: stage timestamps (ns) written by block 0 of the fused thread kernels when
        option ws_dbg has bit 128 set; [64 chunks, 8 stages] int64, zero where nothing was recorded."""
        r = N.lib.mp4j_gpu_rank(self._g) if rank is None else rank
        return self._wrap(r, 114688 + 65536, 64 * 8, N.I64).view(64, 8)

    # raw collectives on lists of tensors (one per local rank)
    def _allreduce(self, ts, code, op, f, t, algo, stream):
        check(N.lib.mp4j_gpu_allreduce(self._g, _ptr_array(ts), code, op, f, t, algo, _stream_ptr(stream)))

    def _allreduce_wire(self, ts, f, t, wire, stream):
        check(N.lib.mp4j_gpu_allreduce_wire(self._g, _ptr_array(ts), f, t, wire, _stream_ptr(stream)))

    def _reduce_scatter(self, ts, code, op, f, counts, algo, stream):
        check(N.lib.mp4j_gpu_reduce_scatter(self._g, _ptr_array(ts), code, op, f, N.i64_array(counts), algo,
                                            _stream_ptr(stream)))

    def _allgather(self, ts, code, froms, tos, algo, stream):
        check(N.lib.mp4j_gpu_allgather(self._g, _ptr_array(ts), code, N.i64_array(froms), N.i64_array(tos), algo,
                                       _stream_ptr(stream)))

    def _broadcast(self, ts, code, f, t, root, algo, stream):
        check(N.lib.mp4j_gpu_broadcast(self._g, _ptr_array(ts), code, f, t, root, algo, _stream_ptr(stream)))

    def _gather(self, ts, code, froms, tos, root, stream):
        check(N.lib.mp4j_gpu_gather(self._g, _ptr_array(ts), code, N.i64_array(froms), N.i64_array(tos), root,
                                    _stream_ptr(stream)))

    def _scatter(self, ts, code, froms, tos, root, stream):
        check(N.lib.mp4j_gpu_scatter(self._g, _ptr_array(ts), code, N.i64_array(froms), N.i64_array(tos), root,
                                     _stream_ptr(stream)))

    def _reduce(self, ts, code, op, f, t, root, algo, stream):
        check(N.lib.mp4j_gpu_reduce(self._g, _ptr_array(ts), code, op, f, t, root, algo, _stream_ptr(stream)))

    def _thread_allreduce(self, ts, T, code, op, f, t, fanout, dst, local_only, thread_algo, algo, stream):
        check(N.lib.mp4j_gpu_thread_allreduce(self._g, _ptr_array(ts), T, code, op, f, t, 1 if fanout else 0, dst,
                                              1 if local_only else 0, thread_algo, algo, _stream_ptr(stream)))


    MAX_SEGMENTS = 128
    SEG_REDUCE_DTYPES = (N.F64, N.F32, N.I64, N.I32, N.I16, N.I8)

    def _thread_segments(self, ts, T, code, op, segs, name, stream):
        """segs: list of (lo, hi, src_rank, src_thread, dst_rank, dst_thread); -1 = all (see mp4j_gpu.h)."""
        flat = [int(x) for s in segs for x in s]
        check(N.lib.mp4j_gpu_thread_segments(self._g, _ptr_array(ts), T, code, op, N.i64_array(flat), len(segs),
                                             name.encode(), _stream_ptr(stream)))

    def _thread_allreduce_host(self, ts, T, code, op, f, t, fanout, dst, stream):
        check(N.lib.mp4j_gpu_thread_allreduce_host(self._g, _ptr_array(ts), T, code, op, f, t, 1 if fanout else 0, dst,
                                                   _stream_ptr(stream)))

    def _lr_xtr_reduce_scatter(self, xts, rts, stream=None):
        """xts[i]: [F, K] bf16, rts[i]: [C, K] bf16 (one pair per local rank). Returns per local rank the
        [fpr, C] fp32 shard (view of plane memory) of G = sum_ranks Xt @ Rt^T, and fpr."""
        F, K = xts[0].shape
        Cc = rts[0].shape[0]
        for x, r in zip(xts, rts):
            if x.dtype != torch.bfloat16 or r.dtype != torch.bfloat16 or not x.is_contiguous() or not r.is_contiguous():
                raise Mp4jException("lr_xtr_reduce_scatter needs contiguous bf16 [F,K] and [C,K] tensors")
            if tuple(x.shape) != (F, K) or tuple(r.shape) != (Cc, K):
                raise Mp4jException("lr_xtr_reduce_scatter: shape mismatch between ranks")
        L = len(xts)
        shards = (C.c_void_p * L)()
        fpr = C.c_int(0)
        check(N.lib.mp4j_gpu_lr_xtr_reduce_scatter(self._g, _ptr_array(xts), _ptr_array(rts), int(F), int(Cc), int(K),
                                                   shards, C.byref(fpr), _stream_ptr(stream)))
        out = []
        for i in range(L):
            t = torch.as_tensor(_RawCuda(shards[i], fpr.value * Cc, N.F32, self), device="cuda:%d" % self.device)
            out.append(t.view(fpr.value, Cc))
        return out, fpr.value

    def _gbdt_hist_reduce_scatter(self, bins, grads, hesss, stream=None):
        """bins[i]: uint8 [F, pitch] feature-major, grads/hesss[i]: fp32 [rows] (one set per local rank).
        Returns per local rank the [fpr, 256, 2] fp32 histogram shard (plane memory) and fpr."""
        F = bins[0].shape[0]
        for b, g, h in zip(bins, grads, hesss):
            if b.dtype != torch.uint8 or b.dim() != 2 or b.stride(1) != 1 or b.shape[0] != F:
                raise Mp4jException("gbdt_hist: bins must be uint8 [F, rows] with unit stride along rows")
            if g.dtype != torch.float32 or h.dtype != torch.float32 or g.numel() != h.numel():
                raise Mp4jException("gbdt_hist: grad/hess must be fp32 of equal length")
        L = len(bins)
        shards = (C.c_void_p * L)()
        fpr = C.c_int(0)
        check(N.lib.mp4j_gpu_gbdt_hist_reduce_scatter(
            self._g, _ptr_array(bins), N.i64_array([b.stride(0) for b in bins]), _ptr_array(grads), _ptr_array(hesss),
            N.i64_array([g.numel() for g in grads]), int(F), shards, C.byref(fpr), _stream_ptr(stream)))
        out = []
        for i in range(L):
            t = torch.as_tensor(_RawCuda(shards[i], fpr.value * 512, N.F32, self), device="cuda:%d" % self.device)
            out.append(t.view(fpr.value, 256, 2))
        return out, fpr.value

    def map_reduce_local(self, keys, vals, op=N.SUM, local_rank=0, stream=None):
        """Reduce-by-key on one GPU (thread-level map merge, DOP:458-477): concatenate the maps of
        the threads of a process into (keys, vals) and get back unique keys with reduced values."""
        _check_tensor(keys)
        _check_tensor(vals)
        if keys.dtype != torch.int64 or keys.numel() != vals.numel():
            raise Mp4jException("map_reduce_local needs int64 keys and as many values as keys")
        n = keys.numel()
        ok = torch.empty(max(n, 1), dtype=torch.int64, device=keys.device)
        ov = torch.empty(max(n, 1), dtype=vals.dtype, device=keys.device)
        cnt = torch.zeros(1, dtype=torch.int64, device=keys.device)
        op = op if isinstance(op, int) else _native_op(op)
        check(N.lib.mp4j_gpu_map_reduce_local(self._g, int(local_rank), keys.data_ptr(), vals.data_ptr(), n,
                                              _code_of(vals), op, ok.data_ptr(), ov.data_ptr(), n, cnt.data_ptr(),
                                              _stream_ptr(stream)))
        m = int(cnt.item())
        return ok[:m], ov[:m]

    def _map_collective(self, mode, keys, vals, max_n, out_cap, op, partition, stream, clone=True):
        """keys/vals: one tensor per local rank. Returns [(keys_out, vals_out, info)] per local rank."""
        L = len(keys)
        for k, v in zip(keys, vals):
            _check_tensor(k)
            _check_tensor(v)
            if k.dtype != torch.int64 or k.numel() != v.numel():
                raise Mp4jException("map collectives need int64 keys and as many values as keys")
        code = _code_of(vals[0])
        ok, ov, info = (C.c_void_p * L)(), (C.c_void_p * L)(), (C.c_void_p * L)()
        check(N.lib.mp4j_gpu_map_collective(self._g, MAP_MODES[mode], _ptr_array(keys), _ptr_array(vals),
                                            N.i64_array([k.numel() for k in keys]), int(max_n), int(out_cap), code, op,
                                            MAP_PARTITIONS[partition], ok, ov, info, _stream_ptr(stream)))
        (stream or torch.cuda.current_stream()).synchronize()
        out = []
        for i in range(L):
            st = (C.c_uint64 * 4).from_address(info[i])
            m, rov, oov, own = int(st[0]), int(st[1]), int(st[2]), int(st[3])
            if rov:
                raise Mp4jException("map collective: a receive region overflowed (very skewed key distribution); "
                                    "raise max_n or pre-aggregate")
            if oov:
                raise Mp4jException("map collective: result has more than out_cap=%d pairs" % out_cap)
            kt = torch.as_tensor(_RawCuda(ok[i], max(m, 1), N.I64, self), device="cuda:%d" % self.device)[:m]
            vt = torch.as_tensor(_RawCuda(ov[i], max(m, 1), code, self), device="cuda:%d" % self.device)[:m]
            out.append((kt.clone(), vt.clone(), own) if clone else (kt, vt, own))
        return out


class GpuPlane(_PlaneBase):
    """GPU data plane of a ProcessCommSlave (one process per GPU)."""

    def __init__(self, comm, device=None, heap_bytes=None, staging_bytes=None):
        if not torch.cuda.is_available():
            raise Mp4jException("the GPU data plane needs a CUDA device")
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", torch.cuda.current_device()))
        self.device = int(device)
        torch.cuda.set_device(self.device)
        torch.cuda.init()
        self._comm = comm
        self._g = C.c_void_p()
        self.tuning = _apply_tuning_table(comm.getSlaveNum())
        check(N.lib.mp4j_gpu_init(comm._h, self.device, int(heap_bytes or 0), int(staging_bytes or 0),
                                  C.byref(self._g)))
        self.thread_algo = THREAD_ALGO[os.environ.get("MP4J_THREAD_ALGO", "auto")]
        comm.debug("gpu plane: " + self.describe(), onlyRank0=False)

    def close(self):
        self._g = None  # owned and destroyed by the native communicator

    def empty(self, n, operand):
        """n-element tensor inside the symmetric heap (collective: every rank must call it in the
        same order). Collectives on it are zero-copy."""
        off = self._alloc_off(int(n) * _ES[operand.code])
        return self._wrap(N.lib.mp4j_gpu_rank(self._g), off, int(n), operand.code)

    def free(self, t):
        """Returns a tensor obtained from empty() to the symmetric heap (collective, like empty(): every rank
        frees in the same order; the tensor must not be used afterwards)."""
        base = N.lib.mp4j_gpu_ptr(self._g, N.lib.mp4j_gpu_rank(self._g), 0)
        check(N.lib.mp4j_gpu_free(self._g, t.data_ptr() - base))

    # --- ProcessCommSlave dispatch targets
    def allreduce(self, t, operand, operator, f, to, algo=N.ALGO_AUTO, stream=None):
        _check_tensor(t)
        if (operand.wire != N.WIRE_LOSSLESS and operand.code == N.F32 and operator.code == N.SUM
                and (t.data_ptr() + 4 * f) % 16 == 0):
            # opt-in lossy wire format (Operands.FLOAT_OPERAND(wire="bf16"|"fp8")): K8
            self._allreduce_wire([t], f, to, operand.wire, stream)
            return
        self._allreduce([t], _code_of(t, operand), self._op(operator, operand.code), f, to, algo, stream)

    def reduce_scatter(self, t, operand, operator, f, counts, algo=N.ALGO_AUTO, stream=None):
        _check_tensor(t)
        self._reduce_scatter([t], _code_of(t, operand), self._op(operator, operand.code), f, counts, algo, stream)

    def allgather(self, t, operand, froms, tos, algo=N.ALGO_AUTO, stream=None):
        _check_tensor(t)
        self._allgather([t], _code_of(t, operand), froms, tos, algo, stream)

    def broadcast(self, t, operand, f, to, root, algo=N.ALGO_AUTO, stream=None):
        _check_tensor(t)
        self._broadcast([t], _code_of(t, operand), f, to, root, algo, stream)

    def gather(self, t, operand, froms, tos, root, stream=None):
        _check_tensor(t)
        self._gather([t], _code_of(t, operand), froms, tos, root, stream)

    def scatter(self, t, operand, froms, tos, root, stream=None):
        _check_tensor(t)
        self._scatter([t], _code_of(t, operand), froms, tos, root, stream)

    def reduce(self, t, operand, operator, f, to, root, algo=N.ALGO_AUTO, stream=None):
        _check_tensor(t)
        self._reduce([t], _code_of(t, operand), self._op(operator, operand.code), f, to, root, algo, stream)

    def lr_xtr_reduce_scatter(self, xt, rt, stream=None):
        """Fused LR gradient + reduce-scatter (K12): returns this rank's [fpr, C] fp32 shard of
        G = sum over ranks of xt @ rt.T (xt: [F, rows] bf16, rt: [C, rows] bf16) and fpr."""
        shards, fpr = self._lr_xtr_reduce_scatter([xt], [rt], stream)
        return shards[0], fpr

    def gbdt_hist_reduce_scatter(self, bins_t, grad, hess, stream=None):
        """Fused GBDT histogram + reduce-scatter (K13): this rank's [fpr, 256, 2] (sum grad, sum hess)
        shard over features [rank*fpr, ...), summed over all ranks, and fpr."""
        shards, fpr = self._gbdt_hist_reduce_scatter([bins_t], [grad], [hess], stream)
        return shards[0], fpr

    # --- sparse maps on (int64 keys, values) tensors
    def _map_sizes(self, n, out_cap, max_n=None):
        """The kernels need two numbers that are the same on every rank: the largest input size and the capacity
        of the result. A caller that knows them (an iterative job with a fixed minibatch) passes them and the call
        involves no host communication at all; otherwise they come from ONE host allgather of the sizes."""
        if max_n is not None and n > max_n:
            raise Mp4jException("map collective: this rank has %d pairs, more than max_n = %d" % (n, max_n))
        if max_n is None or out_cap is None:
            from ..operands import Operands
            p, r = self._comm.getSlaveNum(), self._comm.getRank()
            sizes = np.zeros(p, dtype=np.int64)
            sizes[r] = int(n)
            if p > 1:
                self._comm.allgatherArray(sizes, Operands.LONG_OPERAND(), list(range(p)), list(range(1, p + 1)))
            if max_n is None:
                max_n = int(sizes.max())
            if out_cap is None:
                out_cap = max(1, int(sizes.sum()))
        return int(max_n), int(out_cap)

    def allreduce_map(self, keys, vals, operator, out_cap=None, partition="mix", stream=None, max_n=None, copy=True):
        """allreduceMap on tensors (PCS:2053): every rank gets the merged (keys, values); entries
        with equal keys are reduced with `operator` (SUM/MAX/MIN). Order of the result is
        unspecified (like a HashMap's). out_cap defaults to the sum of all ranks' sizes.
        max_n (largest input size over all ranks) and out_cap, when both given, must be the same on every rank and
        spare the call its host allgather of the sizes; copy=False returns views of plane memory that stay valid until the
        next map collective instead of clones."""
        max_n, out_cap = self._map_sizes(keys.numel(), out_cap, max_n)
        k, v, _ = self._map_collective("allreduce", [keys], [vals], max_n, out_cap, _native_op(operator), partition,
                                       stream, clone=copy)[0]
        return k, v

    def reduce_scatter_map(self, keys, vals, operator, partition="mix", stream=None):
        """reduceScatterMap (PCS:1299) with hash ownership: this rank gets the reduced entries of
        the keys it owns."""
        max_n, out_cap = self._map_sizes(keys.numel(), 1)
        k, v, _ = self._map_collective("reduce_scatter", [keys], [vals], max_n, out_cap, _native_op(operator),
                                       partition, stream)[0]
        return k, v

    def allgather_map(self, keys, vals, out_cap=None, stream=None):
        """allgatherMap (PCS:666): concatenation of all ranks' entries in rank order (no reduction)."""
        max_n, out_cap = self._map_sizes(keys.numel(), out_cap)
        k, v, _ = self._map_collective("allgather", [keys], [vals], max_n, out_cap, N.SUM, "mix", stream)[0]
        return k, v

    # --- ThreadCommSlave dispatch targets
    def thread_allreduce(self, ts, operand, operator, f, to, algo=N.ALGO_AUTO, stream=None):
        for t in ts:
            _check_tensor(t)
        local = self.world == 1
        self._thread_allreduce(ts, len(ts), _code_of(ts[0], operand), self._op(operator, operand.code), f, to, True, 0, local,
                               self.thread_algo if local else 0, algo, stream)

    def thread_allreduce_host(self, ts, operand, operator, f, to, fanout=True, dst=0, stream=None):
        """allreduceArray on HOST tensors (one per thread of this process; one tensor = the process-level
        call): chunked H2D || fused kernel || D2H pipeline inside the library. Pinned tensors run at the
        PCIe rate. The result lands in every tensor (fanout) or only in ts[dst]. Synchronous."""
        for t in ts:
            _check_host_tensor(t)
        self._thread_allreduce_host(ts, len(ts), _code_of(ts[0], operand), self._op(operator, operand.code), f, to,
                                    fanout, dst, stream)

    def thread_segments_ok(self, ts, operand, operator, nseg):
        """Can the one-launch segment kernel run this collective? (else the caller composes it from the
        thread reduce, the dense collective and copies)"""
        if nseg > self.MAX_SEGMENTS or not all(t.is_cuda and t.dim() == 1 and t.is_contiguous() for t in ts):
            return False
        if operator is None:
            return ts[0].dtype in DTYPE_CODES
        return operator.is_native and DTYPE_CODES.get(ts[0].dtype) in self.SEG_REDUCE_DTYPES

    def thread_segments(self, name, ts, operand, operator, segs, stream=None):
        """gather / allgather / broadcast / scatter / reduce-scatter / reduce of the T thread tensors of this
        process (and of every other rank's) in ONE launch; segs as in mp4j_gpu_thread_segments."""
        for t in ts:
            _check_tensor(t)
        op = N.COPY if operator is None else self._op(operator, operand.code)
        self._thread_segments(ts, len(ts), _code_of(ts[0], operand), op, segs, name, stream)

    # --- the same three, called by EVERY thread of a ThreadCommSlave with its own tensor: rendezvous, launch by one
    # thread and error delivery to all T threads are native (mp4j_thread_gpu_*, ThreadComm::delegate); the GIL is
    # released while a thread waits. A thread whose own argument check fails still joins (with no buffer): the other
    # threads get an error instead of waiting for it forever, and it raises its own message afterwards.
    def _thread_native(self, tid, launch_tid, t, operand, operator, host, call):
        err, ptr, code, op = None, 0, 0, N.COPY
        try:
            (_check_host_tensor if host else _check_tensor)(t)
            code = _code_of(t, operand)
            if tid == launch_tid and operator is not None:
                op = self._op(operator, operand.code)
            ptr = t.data_ptr() or 16  # an empty tensor has no address; no element of it is touched
        except Exception as e:  # noqa: BLE001 - surfaced below, after the rendezvous
            err = e
        rc = call(ptr, code, op)
        if err is not None:
            raise err if isinstance(err, Mp4jException) else Mp4jException(repr(err))
        check(rc)

    def thread_native_allreduce(self, tg, tid, t, operand, operator, f, to, host=False, fanout=True, dst=0, stream=None):
        local = self.world == 1
        sp = _stream_ptr(stream)
        if host:
            call = lambda ptr, code, op: N.lib.mp4j_thread_gpu_allreduce_host(  # noqa: E731
                tg, self._g, tid, ptr, code, op, f, to, 1 if fanout else 0, dst, sp)
        else:
            call = lambda ptr, code, op: N.lib.mp4j_thread_gpu_allreduce(  # noqa: E731
                tg, self._g, tid, ptr, code, op, f, to, 1 if fanout else 0, dst, 1 if local else 0,
                self.thread_algo if local else 0, N.ALGO_AUTO, sp)
        self._thread_native(tid, 0, t, operand, operator, host, call)

    def thread_native_segments(self, tg, tid, launch_tid, name, t, operand, operator, segs, stream=None):
        flat = N.i64_array([int(x) for s in segs for x in s])
        sp = _stream_ptr(stream)
        self._thread_native(tid, launch_tid, t, operand, operator, False,
                            lambda ptr, code, op: N.lib.mp4j_thread_gpu_segments(
                                tg, self._g, tid, launch_tid, ptr, code, op, flat, len(segs), name.encode(), sp))

    def thread_reduce(self, ts, operand, operator, f, to, dst, stream=None):
        """Intra-process reduce of the T buffers into ts[dst] (thread path of DOP:441-457)."""
        for t in ts:
            _check_tensor(t)
        self._thread_allreduce(ts, len(ts), _code_of(ts[0], operand), self._op(operator, operand.code), f, to, False, dst, True,
                               0, N.ALGO_AUTO, stream)


class EmuPlane(_PlaneBase):
    """`world` emulated ranks on one GPU. Methods take one tensor per emulated rank."""

    def __init__(self, world, device=0, heap_bytes=256 << 20, staging_bytes=32 << 20):
        self.device = int(device)
        torch.cuda.set_device(self.device)
        torch.cuda.init()
        self._g = C.c_void_p()
        check(N.lib.mp4j_gpu_emu_create(int(world), self.device, int(heap_bytes), int(staging_bytes),
                                        C.byref(self._g)))

    def close(self):
        if self._g:
            torch.cuda.synchronize()
            N.lib.mp4j_gpu_destroy(self._g)
            self._g = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def empty(self, n, code):
        """Symmetric allocation: returns one tensor per emulated rank, all at the same offset."""
        off = self._alloc_off(int(n) * _ES[code])
        return [self._wrap(r, off, int(n), code) for r in range(self.world)]

    def allreduce(self, ts, op, f, t, algo=N.ALGO_AUTO, stream=None):
        self._allreduce(ts, _code_of(ts[0]), op, f, t, algo, stream)

    def allreduce_wire(self, ts, f, t, wire, stream=None):
        self._allreduce_wire(ts, f, t, wire, stream)

    def reduce_scatter(self, ts, op, f, counts, algo=N.ALGO_AUTO, stream=None):
        self._reduce_scatter(ts, _code_of(ts[0]), op, f, counts, algo, stream)

    def allgather(self, ts, froms, tos, algo=N.ALGO_AUTO, stream=None):
        self._allgather(ts, _code_of(ts[0]), froms, tos, algo, stream)

    def broadcast(self, ts, f, t, root, algo=N.ALGO_AUTO, stream=None):
        self._broadcast(ts, _code_of(ts[0]), f, t, root, algo, stream)

    def gather(self, ts, froms, tos, root, stream=None):
        self._gather(ts, _code_of(ts[0]), froms, tos, root, stream)

    def scatter(self, ts, froms, tos, root, stream=None):
        self._scatter(ts, _code_of(ts[0]), froms, tos, root, stream)

    def reduce(self, ts, op, f, t, root, algo=N.ALGO_AUTO, stream=None):
        self._reduce(ts, _code_of(ts[0]), op, f, t, root, algo, stream)

    def lr_xtr_reduce_scatter(self, xts, rts, stream=None):
        return self._lr_xtr_reduce_scatter(xts, rts, stream)

    def gbdt_hist_reduce_scatter(self, bins, grads, hesss, stream=None):
        return self._gbdt_hist_reduce_scatter(bins, grads, hesss, stream)

    def map_collective(self, mode, keys, vals, op=N.SUM, out_cap=None, partition="mix", stream=None):
        """keys/vals: one tensor per emulated rank. Returns [(keys, vals, owned)] per rank."""
        max_n = max(k.numel() for k in keys)
        if out_cap is None:
            out_cap = max(1, sum(k.numel() for k in keys))
        return self._map_collective(mode, keys, vals, max_n, out_cap, op, partition, stream)

    def thread_allreduce(self, ts_by_rank, op, f, t, fanout=True, dst=0, local_only=False, thread_algo=0,
                         algo=N.ALGO_AUTO, stream=None):
        """ts_by_rank: [rank][thread] tensors."""
        T = len(ts_by_rank[0])
        flat = [x for row in ts_by_rank for x in row]
        self._thread_allreduce(flat, T, _code_of(flat[0]), op, f, t, fanout, dst, local_only, thread_algo, algo,
                               stream)

    def thread_segments(self, name, ts_by_rank, op, segs, stream=None):
        """ts_by_rank: [rank][thread] tensors; op: native operator code or N.COPY."""
        T = len(ts_by_rank[0])
        flat = [x for row in ts_by_rank for x in row]
        self._thread_segments(flat, T, _code_of(flat[0]), op, segs, name, stream)

    def thread_allreduce_host(self, ts_by_rank, op, f, t, fanout=True, dst=0, stream=None):
        """ts_by_rank: [rank][thread] HOST tensors (see GpuPlane.thread_allreduce_host)."""
        T = len(ts_by_rank[0])
        flat = [x for row in ts_by_rank for x in row]
        for x in flat:
            _check_host_tensor(x)
        self._thread_allreduce_host(flat, T, _code_of(flat[0]), op, f, t, fanout, dst, stream)
