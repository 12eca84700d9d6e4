package com.fenbi.mp4j.operand;

/** {@code float[]} payloads (and {@code Map<String, Float>} values): mp4j_dtype 1 of the native core. */
public class FloatOperand extends Operand {
    public FloatOperand() { this(false); }
    public FloatOperand(boolean compress) { super("float", 1, compress); }
    @Override public Object newArray(int length) { return new float[length]; }
}
