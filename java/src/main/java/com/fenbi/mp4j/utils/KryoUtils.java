package com.fenbi.mp4j.utils;

import com.fenbi.mp4j.operand.ObjectOperand;

import java.lang.reflect.Method;

/** Optional bridge to Kryo (J/utils/KryoUtils.java:33-46 keeps a thread-local Kryo and hands out default
 *  serializers). The facade itself does not depend on Kryo; when Kryo 4/5 is on the class path this class
 *  builds an {@link ObjectOperand.Codec} that (de)serialises elements with a thread-local Kryo instance,
 *  found by reflection, so user types that are not java.io.Serializable keep working as in the reference. */
public final class KryoUtils {
    private KryoUtils() {}

    private static final ThreadLocal<Object> KRYO = new ThreadLocal<Object>();

    /** The calling thread's Kryo instance (created on first use); throws if Kryo is not available. */
    public static Object getKryo() throws Exception {
        Object k = KRYO.get();
        if (k == null) {
            k = Class.forName("com.esotericsoftware.kryo.Kryo").getDeclaredConstructor().newInstance();
            try {  // Kryo 5 requires registration unless told otherwise
                k.getClass().getMethod("setRegistrationRequired", boolean.class).invoke(k, false);
            } catch (NoSuchMethodException ignored) {
            }
            KRYO.set(k);
        }
        return k;
    }

    /** Codec backed by Kryo's default serializer for `type` (the role of getDefaultSerializer). */
    public static <T> ObjectOperand.Codec<T> getDefaultSerializer(final Class<T> type) {
        return new ObjectOperand.Codec<T>() {
            @Override public byte[] write(T object) throws Exception {
                Object kryo = getKryo();
                Class<?> outCls = Class.forName("com.esotericsoftware.kryo.io.Output");
                Object out = outCls.getConstructor(int.class, int.class).newInstance(256, -1);
                Method w = kryo.getClass().getMethod("writeObject", outCls, Object.class);
                w.invoke(kryo, out, object);
                return (byte[]) outCls.getMethod("toBytes").invoke(out);
            }
            @Override public T read(byte[] bytes) throws Exception {
                Object kryo = getKryo();
                Class<?> inCls = Class.forName("com.esotericsoftware.kryo.io.Input");
                Object in = inCls.getConstructor(byte[].class).newInstance((Object) bytes);
                Method r = kryo.getClass().getMethod("readObject", inCls, Class.class);
                return type.cast(r.invoke(kryo, in, type));
            }
        };
    }
}
