package com.fenbi.mp4j.operator;

/** Reduction over {@code short} values: {@code apply(o1, o2)} (same contract as the reference's IShortOperator).
 *  The predefined constants of {@link Operators.Short} also carry a native code and then run inside the
 *  C++/CUDA core; a user implementation runs on the Java side, element by element. */
public interface IShortOperator extends IOperator {
    short apply(short o1, short o2);
}
