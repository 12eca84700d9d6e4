package com.fenbi.mp4j.operand;

/** {@code long[]} payloads (and {@code Map<String, Long>} values): mp4j_dtype 2 of the native core. */
public class LongOperand extends Operand {
    public LongOperand() { this(false); }
    public LongOperand(boolean compress) { super("long", 2, compress); }
    @Override public Object newArray(int length) { return new long[length]; }
}
