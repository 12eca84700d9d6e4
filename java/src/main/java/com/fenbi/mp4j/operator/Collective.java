package com.fenbi.mp4j.operator;

/** The seven collectives (same constants and order as the reference's enum). */
public enum Collective {
    GATHER("gather"), SCATTER("scatter"), ALL_GATHER("allgather"), BROADCAST("broadcast"),
    REDUCE_SCATTER("reducescatter"), REDUCE("reduce"), ALL_REDUCE("allreduce");

    private final String name;
    Collective(String name) { this.name = name; }
    public String getName() { return name; }
    @Override public String toString() { return name; }
}
