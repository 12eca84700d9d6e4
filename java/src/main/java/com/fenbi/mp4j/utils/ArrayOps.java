package com.fenbi.mp4j.utils;

import com.fenbi.mp4j.exception.Mp4jException;
import com.fenbi.mp4j.operator.IByteOperator;
import com.fenbi.mp4j.operator.IDoubleOperator;
import com.fenbi.mp4j.operator.IFloatOperator;
import com.fenbi.mp4j.operator.IIntOperator;
import com.fenbi.mp4j.operator.ILongOperator;
import com.fenbi.mp4j.operator.IObjectOperator;
import com.fenbi.mp4j.operator.IOperator;
import com.fenbi.mp4j.operator.IShortOperator;
import com.fenbi.mp4j.operator.IStringOperator;

import java.io.ByteArrayInputStream;
import java.io.ByteArrayOutputStream;
import java.io.InputStream;
import java.io.ObjectInputStream;
import java.io.ObjectOutputStream;
import java.io.OutputStream;
import java.lang.reflect.Array;
import java.util.Arrays;
import java.util.Map;
import java.util.zip.DeflaterOutputStream;
import java.util.zip.InflaterInputStream;

/** Type-dispatched array helpers shared by the communicators: element-wise reduce, range copy,
 *  slicing and the byte codec of the generic (Java-side) transport. One switch on the array class
 *  replaces the reference's eight copies of the same loops (DOP:187-205, 362-482 and siblings). */
public final class ArrayOps {
    private ArrayOps() {}

    public static int length(Object arr) { return Array.getLength(arr); }

    /** dst[i] = op(dst[i], src[i + srcShift]) for i in [from, to): the thread-level combine and the
     *  recv-and-reduce step of the generic path, operator order op(local, incoming). */
    @SuppressWarnings("unchecked")
    public static void reduce(Object dst, Object src, int from, int to, int srcShift, IOperator op) throws Mp4jException {
        if (dst instanceof double[]) {
            double[] d = (double[]) dst, s = (double[]) src;
            IDoubleOperator o = (IDoubleOperator) op;
            for (int i = from; i < to; i++) d[i] = o.apply(d[i], s[i + srcShift]);
        } else if (dst instanceof float[]) {
            float[] d = (float[]) dst, s = (float[]) src;
            IFloatOperator o = (IFloatOperator) op;
            for (int i = from; i < to; i++) d[i] = o.apply(d[i], s[i + srcShift]);
        } else if (dst instanceof long[]) {
            long[] d = (long[]) dst, s = (long[]) src;
            ILongOperator o = (ILongOperator) op;
            for (int i = from; i < to; i++) d[i] = o.apply(d[i], s[i + srcShift]);
        } else if (dst instanceof int[]) {
            int[] d = (int[]) dst, s = (int[]) src;
            IIntOperator o = (IIntOperator) op;
            for (int i = from; i < to; i++) d[i] = o.apply(d[i], s[i + srcShift]);
        } else if (dst instanceof short[]) {
            short[] d = (short[]) dst, s = (short[]) src;
            IShortOperator o = (IShortOperator) op;
            for (int i = from; i < to; i++) d[i] = o.apply(d[i], s[i + srcShift]);
        } else if (dst instanceof byte[]) {
            byte[] d = (byte[]) dst, s = (byte[]) src;
            IByteOperator o = (IByteOperator) op;
            for (int i = from; i < to; i++) d[i] = o.apply(d[i], s[i + srcShift]);
        } else if (dst instanceof String[]) {
            String[] d = (String[]) dst, s = (String[]) src;
            IStringOperator o = (IStringOperator) op;
            for (int i = from; i < to; i++) d[i] = o.apply(d[i], s[i + srcShift]);
        } else if (dst instanceof Object[]) {
            Object[] d = (Object[]) dst, s = (Object[]) src;
            IObjectOperator<Object> o = (IObjectOperator<Object>) op;
            for (int i = from; i < to; i++) d[i] = o.apply(d[i], s[i + srcShift]);
        } else {
            throw new Mp4jException("unsupported array type " + dst.getClass());
        }
    }

    /** Applies the operator to two boxed values of the operand's element type (map merges, scalars). */
    @SuppressWarnings("unchecked")
    public static Object apply(IOperator op, Object a, Object b) throws Mp4jException {
        if (op instanceof IDoubleOperator && a instanceof Double) return ((IDoubleOperator) op).apply((Double) a, (Double) b);
        if (op instanceof IFloatOperator && a instanceof Float) return ((IFloatOperator) op).apply((Float) a, (Float) b);
        if (op instanceof ILongOperator && a instanceof Long) return ((ILongOperator) op).apply((Long) a, (Long) b);
        if (op instanceof IIntOperator && a instanceof Integer) return ((IIntOperator) op).apply((Integer) a, (Integer) b);
        if (op instanceof IShortOperator && a instanceof Short) return ((IShortOperator) op).apply((Short) a, (Short) b);
        if (op instanceof IByteOperator && a instanceof Byte) return ((IByteOperator) op).apply((Byte) a, (Byte) b);
        if (op instanceof IStringOperator && a instanceof String) return ((IStringOperator) op).apply((String) a, (String) b);
        if (op instanceof IObjectOperator) return ((IObjectOperator<Object>) op).apply(a, b);
        throw new Mp4jException("operator " + op + " does not match values of " + a.getClass());
    }

    /** local[k] = op(local[k], incoming[k]) or incoming[k] if absent (DOP:225-257). */
    public static <V> void mergeMap(Map<String, V> local, Map<String, V> incoming, IOperator op) throws Mp4jException {
        for (Map.Entry<String, V> e : incoming.entrySet()) {
            V cur = local.get(e.getKey());
            @SuppressWarnings("unchecked")
            V merged = cur == null ? e.getValue() : (V) apply(op, cur, e.getValue());
            local.put(e.getKey(), merged);
        }
    }

    public static void copy(Object src, Object dst, int from, int to) {
        if (to > from && src != dst) System.arraycopy(src, from, dst, from, to - from);
    }

    /** A fresh array holding arr[from, to). */
    public static Object slice(Object arr, int from, int to) {
        if (arr instanceof double[]) return Arrays.copyOfRange((double[]) arr, from, to);
        if (arr instanceof float[]) return Arrays.copyOfRange((float[]) arr, from, to);
        if (arr instanceof long[]) return Arrays.copyOfRange((long[]) arr, from, to);
        if (arr instanceof int[]) return Arrays.copyOfRange((int[]) arr, from, to);
        if (arr instanceof short[]) return Arrays.copyOfRange((short[]) arr, from, to);
        if (arr instanceof byte[]) return Arrays.copyOfRange((byte[]) arr, from, to);
        return Arrays.copyOfRange((Object[]) arr, from, to);
    }

    /** arr[from, from + length(seg)) = seg. */
    public static void put(Object arr, int from, Object seg) {
        int n = length(seg);
        if (n > 0) System.arraycopy(seg, 0, arr, from, n);
    }

    /** One-element array of the operand's element type holding `value` (scalars travel as arrays, PCS:787-829). */
    public static Object box(Object value) {
        if (value instanceof Double) return new double[] {(Double) value};
        if (value instanceof Float) return new float[] {(Float) value};
        if (value instanceof Long) return new long[] {(Long) value};
        if (value instanceof Integer) return new int[] {(Integer) value};
        if (value instanceof Short) return new short[] {(Short) value};
        if (value instanceof Byte) return new byte[] {(Byte) value};
        if (value instanceof String) return new String[] {(String) value};
        return new Object[] {value};
    }

    public static Object unbox(Object arr) { return Array.get(arr, 0); }

    // ---- byte codec of the generic transport: java.io serialization, optionally Deflate-compressed
    // (the reference wraps its Kryo streams in a DeflateSerializer when `compress` is set, DOP:267-269)
    public static byte[] serialize(Object o, boolean compress) throws Mp4jException {
        try {
            ByteArrayOutputStream bos = new ByteArrayOutputStream();
            OutputStream os = compress ? new DeflaterOutputStream(bos) : bos;
            ObjectOutputStream oos = new ObjectOutputStream(os);
            oos.writeObject(o);
            oos.close();
            return bos.toByteArray();
        } catch (Exception e) {
            throw new Mp4jException(e);
        }
    }

    @SuppressWarnings("unchecked")
    public static <X> X deserialize(byte[] b, boolean compress) throws Mp4jException {
        try {
            InputStream is = new ByteArrayInputStream(b);
            if (compress) is = new InflaterInputStream(is);
            ObjectInputStream ois = new ObjectInputStream(is);
            return (X) ois.readObject();
        } catch (Exception e) {
            throw new Mp4jException(e);
        }
    }
}
