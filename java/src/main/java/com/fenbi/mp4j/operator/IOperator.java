package com.fenbi.mp4j.operator;

/** Marker of every operator. Predefined operators carry a native code and run inside the C++/CUDA
 *  core; custom operators implement {@link IObjectOperator} and run on the Java side. */
public interface IOperator {
    /** native enum value (mp4j_op in include/mp4j_types.h) or -1 for a custom operator */
    default int nativeCode() { return -1; }
}
