package com.fenbi.mp4j.operator;

/** Reduction over {@code byte} values: {@code apply(o1, o2)} (same contract as the reference's IByteOperator).
 *  The predefined constants of {@link Operators.Byte} also carry a native code and then run inside the
 *  C++/CUDA core; a user implementation runs on the Java side, element by element. */
public interface IByteOperator extends IOperator {
    byte apply(byte o1, byte o2);
}
