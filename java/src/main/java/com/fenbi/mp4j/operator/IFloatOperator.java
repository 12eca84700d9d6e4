package com.fenbi.mp4j.operator;

/** Reduction over {@code float} values: {@code apply(o1, o2)} (same contract as the reference's IFloatOperator).
 *  The predefined constants of {@link Operators.Float} also carry a native code and then run inside the
 *  C++/CUDA core; a user implementation runs on the Java side, element by element. */
public interface IFloatOperator extends IOperator {
    float apply(float o1, float o2);
}
