package com.fenbi.mp4j.operand;

/** {@code double[]} payloads (and {@code Map<String, Double>} values): mp4j_dtype 0 of the native core. */
public class DoubleOperand extends Operand {
    public DoubleOperand() { this(false); }
    public DoubleOperand(boolean compress) { super("double", 0, compress); }
    @Override public Object newArray(int length) { return new double[length]; }
}
