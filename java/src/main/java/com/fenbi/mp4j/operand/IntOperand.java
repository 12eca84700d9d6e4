package com.fenbi.mp4j.operand;

/** {@code int[]} payloads (and {@code Map<String, Integer>} values): mp4j_dtype 3 of the native core. */
public class IntOperand extends Operand {
    public IntOperand() { this(false); }
    public IntOperand(boolean compress) { super("int", 3, compress); }
    @Override public Object newArray(int length) { return new int[length]; }
}
