package com.fenbi.mp4j.operator;

public enum Collective { GATHER, SCATTER, ALL_GATHER, BROADCAST, REDUCE_SCATTER, REDUCE, ALL_REDUCE }
