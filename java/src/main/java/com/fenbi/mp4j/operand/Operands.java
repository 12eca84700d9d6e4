package com.fenbi.mp4j.operand;

/** Factory with the reference's names (Operands.DOUBLE_OPERAND() ...). */
public final class Operands {
    private Operands() {}
    public static Operand DOUBLE_OPERAND() { return DOUBLE_OPERAND(false); }
    public static Operand FLOAT_OPERAND() { return FLOAT_OPERAND(false); }
    public static Operand LONG_OPERAND() { return LONG_OPERAND(false); }
    public static Operand INT_OPERAND() { return INT_OPERAND(false); }
    public static Operand SHORT_OPERAND() { return SHORT_OPERAND(false); }
    public static Operand BYTE_OPERAND() { return BYTE_OPERAND(false); }
    public static Operand STRING_OPERAND() { return STRING_OPERAND(false); }
    public static Operand OBJECT_OPERAND() { return OBJECT_OPERAND(false); }
    public static Operand DOUBLE_OPERAND(boolean compress) { return new Operand("double", 0, compress); }
    public static Operand FLOAT_OPERAND(boolean compress) { return new Operand("float", 1, compress); }
    public static Operand LONG_OPERAND(boolean compress) { return new Operand("long", 2, compress); }
    public static Operand INT_OPERAND(boolean compress) { return new Operand("int", 3, compress); }
    public static Operand SHORT_OPERAND(boolean compress) { return new Operand("short", 4, compress); }
    public static Operand BYTE_OPERAND(boolean compress) { return new Operand("byte", 5, compress); }
    public static Operand STRING_OPERAND(boolean compress) { return new Operand("string", -1, compress); }
    /** Object payloads travel as java.io.Serializable bytes (the Kryo role of the reference). */
    public static Operand OBJECT_OPERAND(boolean compress) { return new Operand("object", -1, compress); }
}
