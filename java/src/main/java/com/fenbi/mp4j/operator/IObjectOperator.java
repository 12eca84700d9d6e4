package com.fenbi.mp4j.operator;

public interface IObjectOperator<T> extends IOperator {
    T apply(T o1, T o2);
}
