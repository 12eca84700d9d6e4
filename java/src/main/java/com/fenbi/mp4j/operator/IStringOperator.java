package com.fenbi.mp4j.operator;

/** Reduction over {@code String} values: {@code apply(o1, o2)} (same contract as the reference's IStringOperator).
 *  The predefined constants of {@link Operators} also carry a native code and then run inside the
 *  C++/CUDA core; a user implementation runs on the Java side, element by element. */
public interface IStringOperator extends IOperator {
    String apply(String o1, String o2);
}
