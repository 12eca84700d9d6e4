package com.fenbi.mp4j.operand;

/** Factory with the reference's names (J/operand/Operands.java:33-111). */
public final class Operands {
    private Operands() {}
    public static DoubleOperand DOUBLE_OPERAND() { return new DoubleOperand(false); }
    public static DoubleOperand DOUBLE_OPERAND(boolean compress) { return new DoubleOperand(compress); }
    public static FloatOperand FLOAT_OPERAND() { return new FloatOperand(false); }
    public static FloatOperand FLOAT_OPERAND(boolean compress) { return new FloatOperand(compress); }
    public static LongOperand LONG_OPERAND() { return new LongOperand(false); }
    public static LongOperand LONG_OPERAND(boolean compress) { return new LongOperand(compress); }
    public static IntOperand INT_OPERAND() { return new IntOperand(false); }
    public static IntOperand INT_OPERAND(boolean compress) { return new IntOperand(compress); }
    public static ShortOperand SHORT_OPERAND() { return new ShortOperand(false); }
    public static ShortOperand SHORT_OPERAND(boolean compress) { return new ShortOperand(compress); }
    public static ByteOperand BYTE_OPERAND() { return new ByteOperand(false); }
    public static ByteOperand BYTE_OPERAND(boolean compress) { return new ByteOperand(compress); }
    public static StringOperand STRING_OPERAND() { return new StringOperand(false); }
    public static StringOperand STRING_OPERAND(boolean compress) { return new StringOperand(compress); }
    /** Serializable elements. */
    public static <T> ObjectOperand<T> OBJECT_OPERAND(Class<T> type) { return new ObjectOperand<T>(null, type, false); }
    /** Elements with a user codec (the reference passes a Kryo Serializer here). */
    public static <T> ObjectOperand<T> OBJECT_OPERAND(ObjectOperand.Codec<T> codec, Class<T> type) {
        return new ObjectOperand<T>(codec, type, false);
    }
    public static <T> ObjectOperand<T> OBJECT_OPERAND(ObjectOperand.Codec<T> codec, Class<T> type, boolean compress) {
        return new ObjectOperand<T>(codec, type, compress);
    }
}
