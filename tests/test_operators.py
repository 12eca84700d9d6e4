"""Operator unit tests: the value table of the reference's only JUnit class
(src/test/java/com/fenbi/mp4j/operator/OperatorsTest.java:33-117) against the native host
functors (csrc/core/ops.h via mp4j_host_reduce_into) and the Python scalar fallbacks."""
import math

import numpy as np
import pytest

from ytk_mp4j_b200 import Operators, Mp4jException
from ytk_mp4j_b200 import _native as N
from ytk_mp4j_b200.exception import check


def native(dtype_code, np_dtype, op, a, b):
    d = np.array([a], dtype=np_dtype)
    s = np.array([b], dtype=np_dtype)
    check(N.lib.mp4j_host_reduce_into(dtype_code, op.code, d.ctypes.data, s.ctypes.data, 1))
    return d[0]


def test_double_operators():
    D = Operators.Double
    assert native(N.F64, np.float64, D.SUM, 1.5, 2.25) == 3.75
    assert native(N.F64, np.float64, D.MAX, 1.5, 2.25) == 2.25
    assert native(N.F64, np.float64, D.MIN, 1.5, 2.25) == 1.5
    assert native(N.F64, np.float64, D.PROD, 1.5, 2.0) == 3.0
    # Math.max/min semantics: NaN propagates, -0.0 < +0.0
    assert math.isnan(native(N.F64, np.float64, D.MAX, float("nan"), 1.0))
    assert math.isnan(native(N.F64, np.float64, D.MIN, 1.0, float("nan")))
    assert math.copysign(1.0, native(N.F64, np.float64, D.MAX, -0.0, 0.0)) == 1.0
    assert math.copysign(1.0, native(N.F64, np.float64, D.MIN, 0.0, -0.0)) == -1.0


def test_double_loc_operators():
    D = Operators.Double
    a = D.compositeDouble(3.5, 7)
    b = D.compositeDouble(4.5, 9)
    assert D.getFloatVal(a) == 3.5 and D.getIntLoc(a) == 7
    assert D.getFloatVal(b) == 4.5 and D.getIntLoc(b) == 9
    mx = native(N.F64, np.float64, D.FLOAT_MAX_LOC, a, b)
    mn = native(N.F64, np.float64, D.FLOAT_MIN_LOC, a, b)
    assert D.getIntLoc(mx) == 9 and D.getFloatVal(mx) == 4.5
    assert D.getIntLoc(mn) == 7 and D.getFloatVal(mn) == 3.5
    # deterministic tie-break: lower loc wins, either argument order
    t1, t2 = D.compositeDouble(1.0, 5), D.compositeDouble(1.0, 2)
    assert D.getIntLoc(native(N.F64, np.float64, D.FLOAT_MAX_LOC, t1, t2)) == 2
    assert D.getIntLoc(native(N.F64, np.float64, D.FLOAT_MAX_LOC, t2, t1)) == 2
    assert D.getIntLoc(D.FLOAT_MAX_LOC.apply(t1, t2)) == 2
    # negative locations survive the packing
    c = D.compositeDouble(-2.0, -3)
    assert D.getFloatVal(c) == -2.0 and D.getIntLoc(c) == -3


def test_float_operators():
    F = Operators.Float
    assert native(N.F32, np.float32, F.SUM, 1.5, 2.25) == np.float32(3.75)
    assert native(N.F32, np.float32, F.MAX, -1.0, 2.0) == 2.0
    assert native(N.F32, np.float32, F.MIN, -1.0, 2.0) == -1.0
    assert native(N.F32, np.float32, F.PROD, 3.0, 2.0) == 6.0
    assert math.isnan(native(N.F32, np.float32, F.MAX, float("nan"), 1.0))


def test_long_operators():
    L = Operators.Long
    assert native(N.I64, np.int64, L.SUM, 1 << 40, 5) == (1 << 40) + 5
    assert native(N.I64, np.int64, L.MAX, -3, 5) == 5
    assert native(N.I64, np.int64, L.MIN, -3, 5) == -3
    assert native(N.I64, np.int64, L.PROD, -3, 5) == -15
    assert native(N.I64, np.int64, L.BITS_AND, 0b1100, 0b1010) == 0b1000
    assert native(N.I64, np.int64, L.BITS_OR, 0b1100, 0b1010) == 0b1110
    assert native(N.I64, np.int64, L.BITS_XOR, 0b1100, 0b1010) == 0b0110
    a, b = L.compositeLong(10, 1), L.compositeLong(-20, 2)
    assert L.getIntVal(b) == -20 and L.getIntLoc(b) == 2
    assert L.getIntLoc(native(N.I64, np.int64, L.INT_MAX_LOC, a, b)) == 1
    assert L.getIntLoc(native(N.I64, np.int64, L.INT_MIN_LOC, a, b)) == 2
    assert L.getIntLoc(L.INT_MIN_LOC.apply(a, b)) == 2


@pytest.mark.parametrize("grp,code,dt", [(Operators.Int, N.I32, np.int32), (Operators.Short, N.I16, np.int16),
                                         (Operators.Byte, N.I8, np.int8)])
def test_int_short_byte_operators(grp, code, dt):
    assert native(code, dt, grp.SUM, 3, 4) == 7
    assert native(code, dt, grp.MAX, 3, -4) == 3
    assert native(code, dt, grp.MIN, 3, -4) == -4
    assert native(code, dt, grp.PROD, 3, -4) == -12
    assert native(code, dt, grp.BITS_AND, 6, 3) == 2
    assert native(code, dt, grp.BITS_OR, 6, 3) == 7
    assert native(code, dt, grp.BITS_XOR, 6, 3) == 5
    # narrowing wrap-around
    info = np.iinfo(dt)
    assert native(code, dt, grp.SUM, info.max, 1) == info.min


def test_invalid_combinations_raise():
    d = np.zeros(1, np.float32)
    assert N.lib.mp4j_host_reduce_into(N.F32, N.BAND, d.ctypes.data, d.ctypes.data, 1) != 0
    with pytest.raises(Mp4jException):
        check(N.lib.mp4j_host_reduce_into(N.I32, N.MAXLOC, d.ctypes.data, d.ctypes.data, 1))


def test_java_string_hash():
    from ytk_mp4j_b200.comm.process import java_string_hash, key_partition
    assert java_string_hash("") == 0
    assert java_string_hash("hello") == 99162322
    assert java_string_hash("-1") == 1444
    assert java_string_hash("polygenelubricants") == -2147483648  # famous Integer.MIN_VALUE hash
    assert 0 <= key_partition("polygenelubricants", 3) < 3
    assert java_string_hash("é中") == (0xe9 * 31 + 0x4e2d)
