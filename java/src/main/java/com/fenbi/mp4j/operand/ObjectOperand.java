package com.fenbi.mp4j.operand;

import com.fenbi.mp4j.exception.Mp4jException;

import java.io.Serializable;
import java.lang.reflect.Array;
import java.util.ArrayList;
import java.util.HashMap;
import java.util.List;
import java.util.Map;

/** {@code T[]} payloads and {@code Map<String, T>} values of a user type (reduce with IObjectOperator).
 *
 *  The reference takes a Kryo {@code Serializer<T>}; this facade does not depend on Kryo: elements must be
 *  java.io.Serializable, or the caller passes a {@link Codec} that turns one element into bytes and back
 *  (which is also where a Kryo serializer can be plugged in, see KryoUtils). */
public class ObjectOperand<T> extends Operand {
    /** Element codec for types that are not Serializable (the role of the reference's Serializer<T>). */
    public interface Codec<T> {
        byte[] write(T object) throws Exception;
        T read(byte[] bytes) throws Exception;
    }

    private final Class<T> type;
    private final Codec<T> codec;

    public ObjectOperand(Class<T> type) { this(null, type, false); }
    public ObjectOperand(Codec<T> codec, Class<T> type) { this(codec, type, false); }
    public ObjectOperand(Codec<T> codec, Class<T> type, boolean compress) {
        super("object", -1, compress);
        this.type = type;
        this.codec = codec;
    }

    public Class<T> getType() { return type; }
    public Codec<T> getCodec() { return codec; }
    /** name of the reference's accessor for its Kryo serializer */
    public Codec<T> getSerializer() { return codec; }

    @Override public Object newArray(int length) {
        return Array.newInstance(type == null ? Object.class : type, length);
    }

    /** One codec-encoded element inside a serialized payload. */
    private static final class Enc implements Serializable {
        private static final long serialVersionUID = 1L;
        final byte[] bytes;
        Enc(byte[] bytes) { this.bytes = bytes; }
    }

    // With a codec, payloads (element arrays, maps, lists and maps of those) are walked and every
    // element is replaced by its codec bytes before java.io serialization, and restored afterwards.
    @SuppressWarnings("unchecked")
    private Object encode(Object p) throws Exception {
        if (p == null) return null;
        if (type != null && type.isInstance(p)) return new Enc(codec.write((T) p));
        if (p instanceof Object[]) {
            Object[] a = (Object[]) p, out = new Object[a.length];
            for (int i = 0; i < a.length; i++) out[i] = encode(a[i]);
            return out;
        }
        if (p instanceof Map) {
            Map<Object, Object> out = new HashMap<Object, Object>();
            for (Map.Entry<Object, Object> e : ((Map<Object, Object>) p).entrySet()) out.put(e.getKey(), encode(e.getValue()));
            return out;
        }
        if (p instanceof List) {
            List<Object> out = new ArrayList<Object>();
            for (Object o : (List<Object>) p) out.add(encode(o));
            return out;
        }
        return p;
    }

    @SuppressWarnings("unchecked")
    private Object decode(Object p, boolean elementArray) throws Exception {
        if (p == null) return null;
        if (p instanceof Enc) return codec.read(((Enc) p).bytes);
        if (p instanceof Object[]) {
            Object[] a = (Object[]) p;
            boolean elems = a.length > 0 && a[0] instanceof Enc;
            Object[] out = elems ? (Object[]) newArray(a.length) : new Object[a.length];
            for (int i = 0; i < a.length; i++) out[i] = decode(a[i], false);
            return out;
        }
        if (p instanceof Map) {
            Map<Object, Object> out = new HashMap<Object, Object>();
            for (Map.Entry<Object, Object> e : ((Map<Object, Object>) p).entrySet()) out.put(e.getKey(), decode(e.getValue(), false));
            return out;
        }
        if (p instanceof List) {
            List<Object> out = new ArrayList<Object>();
            for (Object o : (List<Object>) p) out.add(decode(o, false));
            return out;
        }
        return p;
    }

    @Override public byte[] toBytes(Object payload) throws Mp4jException {
        if (codec == null) return super.toBytes(payload);
        try {
            return super.toBytes(encode(payload));
        } catch (Mp4jException e) {
            throw e;
        } catch (Exception e) {
            throw new Mp4jException(e);
        }
    }

    @SuppressWarnings("unchecked")
    @Override public <X> X fromBytes(byte[] bytes) throws Mp4jException {
        if (codec == null) return super.<X>fromBytes(bytes);
        try {
            return (X) decode(super.<Object>fromBytes(bytes), false);
        } catch (Mp4jException e) {
            throw e;
        } catch (Exception e) {
            throw new Mp4jException(e);
        }
    }

    /** Bytes of one element for the RPC path (ObjectOperand.convertToBytes of the reference). */
    public byte[] convertToBytes(T object) throws Mp4jException {
        try {
            return codec != null ? codec.write(object) : com.fenbi.mp4j.utils.ArrayOps.serialize(object, false);
        } catch (Mp4jException e) {
            throw e;
        } catch (Exception e) {
            throw new Mp4jException(e);
        }
    }

    public T readFromBytes(byte[] bytes) throws Mp4jException {
        try {
            if (codec != null) return codec.read(bytes);
            return com.fenbi.mp4j.utils.ArrayOps.<T>deserialize(bytes, false);
        } catch (Mp4jException e) {
            throw e;
        } catch (Exception e) {
            throw new Mp4jException(e);
        }
    }
}
