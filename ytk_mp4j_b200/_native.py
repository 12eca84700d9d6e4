"""ctypes binding of the C ABI in include/mp4j.h (libmp4j.so, built in-tree by build.py).

The library is the product: there is no Python fallback for anything it implements. If it is
missing this module builds it once (nvcc cross-compiles without a GPU); if that fails the import
fails loudly.
"""
import ctypes as C
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_lib", "libmp4j.so")

# enums of include/mp4j_types.h
F64, F32, I64, I32, I16, I8, BF16, F16 = range(8)
SUM, MAX, MIN, PROD, BAND, BOR, BXOR, MAXLOC, MINLOC = range(9)
COPY = 15
LOG_DEBUG, LOG_INFO, LOG_ERROR = 0, 1, 2
WIRE_LOSSLESS, WIRE_BF16, WIRE_FP8 = 0, 1, 2
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS = 0, 1, 2, 3
GATHER, SCATTER, ALL_GATHER, BROADCAST, REDUCE_SCATTER, REDUCE, ALL_REDUCE = range(7)  # mp4j_collective
MAX_RANKS = 8
MAX_THREADS = 64


def _load():
    if not os.path.exists(_LIB_PATH):
        sys.path.insert(0, os.path.dirname(_HERE))
        try:
            import build as _build  # repo-root build.py
            _build.build()
        finally:
            sys.path.pop(0)
    return C.CDLL(_LIB_PATH, mode=C.RTLD_GLOBAL)


lib = _load()

c_i64p = C.POINTER(C.c_int64)
c_vpp = C.POINTER(C.c_void_p)


def _proto(name, restype, *argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = list(argtypes)
    return fn


_proto("mp4j_last_error", C.c_char_p)
_proto("mp4j_version", C.c_char_p)
_proto("mp4j_free", None, C.c_void_p)
# master
_proto("mp4j_master_create", C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_char_p,
       C.c_char_p, C.c_int, c_vpp)
_proto("mp4j_master_start", C.c_int, C.c_void_p)
_proto("mp4j_master_port", C.c_int, C.c_void_p)
_proto("mp4j_master_wait", C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int))
_proto("mp4j_master_shutdown", C.c_int, C.c_void_p)
_proto("mp4j_master_logs", C.c_int, C.c_void_p, c_vpp, C.POINTER(C.c_size_t))
_proto("mp4j_master_destroy", None, C.c_void_p)
# comm
_proto("mp4j_comm_create", C.c_int, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_char_p, c_vpp)
_proto("mp4j_comm_destroy", None, C.c_void_p)
_proto("mp4j_comm_rank", C.c_int, C.c_void_p)
_proto("mp4j_comm_size", C.c_int, C.c_void_p)
_proto("mp4j_comm_host_transport", C.c_char_p, C.c_void_p)
_proto("mp4j_host_map_reduce", C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_size_t, C.c_void_p, C.c_size_t,
       C.c_int, C.c_int, c_vpp, C.POINTER(C.c_size_t), c_vpp, C.POINTER(C.c_size_t))
_proto("mp4j_comm_barrier", C.c_int, C.c_void_p)
_proto("mp4j_comm_log", C.c_int, C.c_void_p, C.c_int, C.c_char_p)
_proto("mp4j_comm_close", C.c_int, C.c_void_p, C.c_int)
_proto("mp4j_comm_exchange", C.c_int, C.c_void_p, C.POINTER(C.c_int))
_proto("mp4j_comm_write_file", C.c_int, C.c_void_p, C.c_char_p, C.c_char_p)
_proto("mp4j_comm_shutdown_master", C.c_int, C.c_void_p)
_proto("mp4j_comm_gather_all", C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, c_vpp, C.POINTER(C.c_size_t))
_proto("mp4j_host_reduce_scatter", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64, c_i64p)
_proto("mp4j_host_allgather", C.c_int, C.c_void_p, C.c_void_p, C.c_int, c_i64p, c_i64p)
_proto("mp4j_host_gather", C.c_int, C.c_void_p, C.c_void_p, C.c_int, c_i64p, c_i64p, C.c_int)
_proto("mp4j_host_scatter", C.c_int, C.c_void_p, C.c_void_p, C.c_int, c_i64p, c_i64p, C.c_int)
_proto("mp4j_host_broadcast", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int)
_proto("mp4j_host_reduce", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_int64, C.c_int)
_proto("mp4j_host_allreduce", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64, C.c_int64)
_proto("mp4j_host_allreduce_rpc", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int64)
_proto("mp4j_host_reduce_into", C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int64)
_proto("mp4j_comm_send_bytes", C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t)
_proto("mp4j_comm_recv_bytes", C.c_int, C.c_void_p, C.c_int, c_vpp, C.POINTER(C.c_size_t))
_proto("mp4j_comm_sendrecv_bytes", C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_int, c_vpp,
       C.POINTER(C.c_size_t))
# thread group
_proto("mp4j_tgroup_create", C.c_int, C.c_int, c_vpp)
_proto("mp4j_tgroup_destroy", None, C.c_void_p)
_proto("mp4j_tgroup_barrier", C.c_int, C.c_void_p)
_proto("mp4j_tgroup_set_slot", C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_uint64)
_proto("mp4j_tgroup_get_slot", C.c_uint64, C.c_void_p, C.c_int, C.c_int)
_proto("mp4j_thread_collective", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
       C.c_int64, C.c_int64, c_i64p, c_i64p, c_i64p, C.c_int, C.c_int)
_proto("mp4j_thread_allreduce", C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int64,
       C.c_int64)
# helpers
_proto("mp4j_composite_double", C.c_double, C.c_float, C.c_int32)
_proto("mp4j_composite_double_val", C.c_float, C.c_double)
_proto("mp4j_composite_double_loc", C.c_int32, C.c_double)
_proto("mp4j_composite_long", C.c_int64, C.c_int32, C.c_int32)
_proto("mp4j_composite_long_val", C.c_int32, C.c_int64)
_proto("mp4j_composite_long_loc", C.c_int32, C.c_int64)
_proto("mp4j_java_string_hash", C.c_int32, C.c_char_p, C.c_size_t)
_proto("mp4j_host_map_reduce_scatter", C.c_int, C.c_void_p, C.c_char_p, C.c_size_t, C.c_void_p, c_i64p, C.c_int, C.c_int,
       c_vpp, C.POINTER(C.c_size_t), c_vpp, C.POINTER(C.c_size_t))
BLOB_GET = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t))
BLOB_PUT = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t)
_proto("mp4j_host_blob_collective", C.c_int, C.c_void_p, C.c_int, C.c_int, BLOB_GET, BLOB_PUT, C.c_void_p)
_proto("mp4j_map_keys64", C.c_int, C.c_char_p, C.c_size_t, C.c_int64, c_i64p)


def last_error():
    return lib.mp4j_last_error().decode("utf-8", "replace")


def i64_array(values):
    arr = (C.c_int64 * len(values))(*[int(v) for v in values])
    return arr


def take_bytes(ptr, length):
    """Copies a malloc'd (ptr, len) result into bytes and frees it."""
    try:
        return C.string_at(ptr, length.value)
    finally:
        lib.mp4j_free(ptr)
