package com.fenbi.mp4j.operator;

/** Reduction over {@code int} values: {@code apply(o1, o2)} (same contract as the reference's IIntOperator).
 *  The predefined constants of {@link Operators.Int} also carry a native code and then run inside the
 *  C++/CUDA core; a user implementation runs on the Java side, element by element. */
public interface IIntOperator extends IOperator {
    int apply(int o1, int o2);
}
