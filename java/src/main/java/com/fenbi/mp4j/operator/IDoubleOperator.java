package com.fenbi.mp4j.operator;

/** Reduction over {@code double} values: {@code apply(o1, o2)} (same contract as the reference's IDoubleOperator).
 *  The predefined constants of {@link Operators.Double} also carry a native code and then run inside the
 *  C++/CUDA core; a user implementation runs on the Java side, element by element. */
public interface IDoubleOperator extends IOperator {
    double apply(double o1, double o2);
}
