"""Collective / Container enums (parity: J/operator/Collective.java:29-36, Container.java:29-31)."""
import enum


class Collective(enum.IntEnum):
    GATHER = 0
    SCATTER = 1
    ALL_GATHER = 2
    BROADCAST = 3
    REDUCE_SCATTER = 4
    REDUCE = 5
    ALL_REDUCE = 6


class Container(enum.IntEnum):
    ARRAY = 0
    MAP = 1
