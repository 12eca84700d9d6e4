package com.fenbi.mp4j.operand;

/** {@code String[]} payloads and {@code Map<String, String>} values; no predefined operators exist for
 *  strings (implement IStringOperator). Always takes the Java-side generic path (host TCP transport). */
public class StringOperand extends Operand {
    public StringOperand() { this(false); }
    public StringOperand(boolean compress) { super("string", -1, compress); }
    @Override public Object newArray(int length) { return new String[length]; }
}
