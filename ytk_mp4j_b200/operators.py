"""Operators: predefined reductions and custom operators.

Parity: J/operator/Operators.java:30-351 (Double/Float/Long/Int/Short/Byte groups, packed LOC
helpers) and the I<T>Operator interfaces (J/operator/I*Operator.java:29-30).

A predefined operator is a native op code: it runs inside the CUDA kernels (device functors of
csrc/core/ops.h) or inside the C++ host collectives. A custom operator wraps a Python callable
`apply(a, b)`; it runs on the host path (element-wise, like the reference's interface call).
"""
from . import _native as N


class Operator:
    """A reduction operator. `code` is the native enum or None for custom operators."""

    def __init__(self, name, code=None, fn=None, vectorized=False, expr=None):
        self.name = name
        self.code = code
        self.fn = fn
        self.vectorized = vectorized
        self.expr = expr  # C++ expression over a, b: compiled for the GPU path (device_expr)

    def apply(self, a, b):
        if self.fn is not None:
            return self.fn(a, b)
        return _PY_FALLBACK[self.code](a, b)

    __call__ = apply

    @property
    def is_native(self):
        return self.code is not None

    def __repr__(self):
        return "Operator(%s)" % self.name


def custom(fn, name="custom", vectorized=False):
    """Custom operator from a Python callable (I<T>Operator.apply). With vectorized=True the
    callable receives whole numpy segments instead of scalars."""
    return Operator(name, None, fn, vectorized)


def device_expr(expr, name=None, host_fn=None, vectorized=False):
    """Custom operator for CUDA tensors: `expr` is a C++ expression over `a` (accumulated so far, in
    ascending rank / thread order) and `b` (the next value), both of the operand's element type, e.g.
    "a + 2 * b" or "fabs(a) > fabs(b) ? a : b". It is compiled once per (element type, world size) with
    NVRTC into the same fused peer-memory kernels as the predefined operators
    (csrc/gpu/custom_op.cu) - the GPU form of implementing I<T>Operator.apply. `host_fn(a, b)` is the
    Python equivalent used when the same operator meets numpy arrays / lists on the host path."""
    def _no_host(a, b):
        from .exception import Mp4jException
        raise Mp4jException("operator %r was given as a device expression without host_fn: "
                            "it only applies to CUDA tensors" % (name or expr))
    return Operator(name or "expr(%s)" % expr, None, host_fn or _no_host, vectorized, expr)


def _maxloc_d(a, b):
    va, vb = Double.getFloatVal(a), Double.getFloatVal(b)
    if va != vb:
        return a if va > vb else b
    return a if Double.getIntLoc(a) <= Double.getIntLoc(b) else b


def _minloc_d(a, b):
    va, vb = Double.getFloatVal(a), Double.getFloatVal(b)
    if va != vb:
        return a if va < vb else b
    return a if Double.getIntLoc(a) <= Double.getIntLoc(b) else b


def _maxloc_l(a, b):
    va, vb = Long.getIntVal(a), Long.getIntVal(b)
    if va != vb:
        return a if va > vb else b
    return a if Long.getIntLoc(a) <= Long.getIntLoc(b) else b


def _minloc_l(a, b):
    va, vb = Long.getIntVal(a), Long.getIntVal(b)
    if va != vb:
        return a if va < vb else b
    return a if Long.getIntLoc(a) <= Long.getIntLoc(b) else b


# scalar fallbacks (used for map values and scalars on the Python side)
_PY_FALLBACK = {
    N.SUM: lambda a, b: a + b,
    N.MAX: lambda a, b: a if a >= b else b,
    N.MIN: lambda a, b: a if a <= b else b,
    N.PROD: lambda a, b: a * b,
    N.BAND: lambda a, b: a & b,
    N.BOR: lambda a, b: a | b,
    N.BXOR: lambda a, b: a ^ b,
}


class _Group:
    SUM = Operator("SUM", N.SUM)
    MAX = Operator("MAX", N.MAX)
    MIN = Operator("MIN", N.MIN)
    PROD = Operator("PROD", N.PROD)


class _IntGroup(_Group):
    BITS_AND = Operator("BITS_AND", N.BAND)
    BITS_OR = Operator("BITS_OR", N.BOR)
    BITS_XOR = Operator("BITS_XOR", N.BXOR)


class Double(_Group):
    """Operators.Double (Operators.java:30-89)."""
    FLOAT_MAX_LOC = Operator("FLOAT_MAX_LOC", N.MAXLOC, _maxloc_d)
    FLOAT_MIN_LOC = Operator("FLOAT_MIN_LOC", N.MINLOC, _minloc_d)

    @staticmethod
    def compositeDouble(val, loc):
        return N.lib.mp4j_composite_double(float(val), int(loc))

    @staticmethod
    def getFloatVal(composite):
        return N.lib.mp4j_composite_double_val(float(composite))

    @staticmethod
    def getIntLoc(composite):
        return N.lib.mp4j_composite_double_loc(float(composite))


class Float(_Group):
    """Operators.Float (Operators.java:91-117)."""


class Long(_IntGroup):
    """Operators.Long (Operators.java:119-202)."""
    INT_MAX_LOC = Operator("INT_MAX_LOC", N.MAXLOC, _maxloc_l)
    INT_MIN_LOC = Operator("INT_MIN_LOC", N.MINLOC, _minloc_l)

    @staticmethod
    def compositeLong(val, loc):
        return N.lib.mp4j_composite_long(int(val), int(loc))

    @staticmethod
    def getIntVal(composite):
        return N.lib.mp4j_composite_long_val(int(composite))

    @staticmethod
    def getIntLoc(composite):
        return N.lib.mp4j_composite_long_loc(int(composite))


class Int(_IntGroup):
    """Operators.Int (Operators.java:204-252)."""


class Short(_IntGroup):
    """Operators.Short (Operators.java:254-302)."""


class Byte(_IntGroup):
    """Operators.Byte (Operators.java:304-351)."""


class Operators:
    """Namespace mirror of the reference's `Operators` class."""
    Double = Double
    Float = Float
    Long = Long
    Int = Int
    Short = Short
    Byte = Byte
    custom = staticmethod(custom)
    device_expr = staticmethod(device_expr)
