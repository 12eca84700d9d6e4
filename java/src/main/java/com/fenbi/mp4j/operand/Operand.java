package com.fenbi.mp4j.operand;

/** Typed payload descriptor. Immutable (the reference's operands are stateful per call). */
public final class Operand {
    /** mp4j_dtype of include/mp4j_types.h, or -1 for String/Object payloads */
    public final int dtype;
    public final boolean compress;
    public final String name;

    Operand(String name, int dtype, boolean compress) {
        this.name = name;
        this.dtype = dtype;
        this.compress = compress;
    }
    public boolean isPrimitive() { return dtype >= 0; }
}
