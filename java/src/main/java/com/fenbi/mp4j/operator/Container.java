package com.fenbi.mp4j.operator;

/** Payload container kinds. */
public enum Container {
    ARRAY("array"), MAP("map");

    private final String name;
    Container(String name) { this.name = name; }
    public String getName() { return name; }
    @Override public String toString() { return name; }
}
