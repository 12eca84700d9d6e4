package com.fenbi.mp4j.operator;

/** Reduction over {@code long} values: {@code apply(o1, o2)} (same contract as the reference's ILongOperator).
 *  The predefined constants of {@link Operators.Long} also carry a native code and then run inside the
 *  C++/CUDA core; a user implementation runs on the Java side, element by element. */
public interface ILongOperator extends IOperator {
    long apply(long o1, long o2);
}
