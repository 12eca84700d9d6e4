"""Operands: typed payload descriptors (parity: J/operand/Operands.java:33-111, Operand.java:42-80).

An operand names the element type, whether the wire should be compressed, and - for
String/Object - how elements are (de)serialised. Unlike the reference's operands
(`collective`, `container`, `operator` fields swapped per call, PCS:552-554) these objects are
immutable and can be shared between threads and calls.

compress=True:
  * host path (numpy / String / Object / maps): zlib Deflate over the serialised message, i.e. the
    reference's DeflateSerializer behaviour (DOP:267-269), lossless;
  * GPU path: ignored unless `wire` is set, because the GPU wire formats ("bf16", "fp8" block
    scaled) are lossy for float/double; lossless stays the default.
"""
import pickle
import zlib

import numpy as np

from . import _native as N
from .exception import Mp4jException

_WIRE = {None: N.WIRE_LOSSLESS, "lossless": N.WIRE_LOSSLESS, "bf16": N.WIRE_BF16, "fp8": N.WIRE_FP8}


class PickleSerializer:
    """Default element codec for String/Object payloads (the Kryo role, KryoUtils.java:33-46)."""

    def dumps(self, obj):
        return pickle.dumps(obj, protocol=pickle.HIGHEST_PROTOCOL)

    def loads(self, data):
        return pickle.loads(data)


class Operand:
    def __init__(self, name, code, np_dtype, compress=False, wire=None, serializer=None, py_type=None):
        if wire not in _WIRE:
            raise Mp4jException("unknown wire format %r (use None, 'bf16' or 'fp8')" % (wire,))
        self.name = name
        self.code = code            # native dtype enum, None for String/Object
        self.np_dtype = np_dtype
        self.compress = bool(compress)
        self.wire = _WIRE[wire]
        self.serializer = serializer or PickleSerializer()
        self.py_type = py_type

    @property
    def is_primitive(self):
        return self.code is not None

    # message codec of the host generic path
    def dumps(self, payload):
        data = self.serializer.dumps(payload)
        return zlib.compress(data, 1) if self.compress else data

    def loads(self, data):
        if self.compress:
            data = zlib.decompress(data)
        return self.serializer.loads(data)

    def cast(self, value):
        """Coerces a python scalar to the operand's element type (wrap-around for ints)."""
        if self.np_dtype is None:
            return value
        return np.array(value).astype(self.np_dtype, casting="unsafe").item()

    def __repr__(self):
        return "Operand(%s%s)" % (self.name, ", compress" if self.compress else "")


class Operands:
    """Factory mirror of the reference's `Operands` class."""

    @staticmethod
    def DOUBLE_OPERAND(compress=False, wire=None):
        return Operand("double", N.F64, np.float64, compress, wire)

    @staticmethod
    def FLOAT_OPERAND(compress=False, wire=None):
        return Operand("float", N.F32, np.float32, compress, wire)

    @staticmethod
    def LONG_OPERAND(compress=False):
        return Operand("long", N.I64, np.int64, compress)

    @staticmethod
    def INT_OPERAND(compress=False):
        return Operand("int", N.I32, np.int32, compress)

    @staticmethod
    def SHORT_OPERAND(compress=False):
        return Operand("short", N.I16, np.int16, compress)

    @staticmethod
    def BYTE_OPERAND(compress=False):
        return Operand("byte", N.I8, np.int8, compress)

    @staticmethod
    def STRING_OPERAND(compress=False):
        return Operand("string", None, None, compress, py_type=str)

    @staticmethod
    def OBJECT_OPERAND(serializer=None, type=None, compress=False):
        return Operand("object", None, None, compress, serializer=serializer, py_type=type)

    # additions for torch users (no counterpart in the reference)
    @staticmethod
    def BF16_OPERAND():
        return Operand("bf16", N.BF16, None)

    @staticmethod
    def HALF_OPERAND():
        return Operand("half", N.F16, np.float16)
