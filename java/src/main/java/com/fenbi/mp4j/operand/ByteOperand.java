package com.fenbi.mp4j.operand;

/** {@code byte[]} payloads (and {@code Map<String, Byte>} values): mp4j_dtype 5 of the native core. */
public class ByteOperand extends Operand {
    public ByteOperand() { this(false); }
    public ByteOperand(boolean compress) { super("byte", 5, compress); }
    @Override public Object newArray(int length) { return new byte[length]; }
}
