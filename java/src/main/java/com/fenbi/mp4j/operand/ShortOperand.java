package com.fenbi.mp4j.operand;

/** {@code short[]} payloads (and {@code Map<String, Short>} values): mp4j_dtype 4 of the native core. */
public class ShortOperand extends Operand {
    public ShortOperand() { this(false); }
    public ShortOperand(boolean compress) { super("short", 4, compress); }
    @Override public Object newArray(int length) { return new short[length]; }
}
