package com.fenbi.mp4j.operator;

public enum Container { ARRAY, MAP }
