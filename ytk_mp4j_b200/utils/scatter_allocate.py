"""Scatter-tree planner (parity in purpose with J/utils/ScatterAllocate.java:58-131).

The GPU path needs no tree (every peer pulls its slice from the root over NVSwitch in one
kernel), and the C++ host plane uses a binomial tree (csrc/host/process_comm.cc). This planner
states that binomial schedule explicitly so tests can assert the reference's invariant
(`recvNum`: every non-root rank receives exactly one message, the root none).
"""


def allocate(slave_num, root):
    """Returns {src_rank: [(src, dst, rank_from, rank_to_inclusive), ...]} in send order.
    rank_from/rank_to are RELATIVE ranks (v = (rank - root) mod p) mapped back to real ranks."""
    p = slave_num
    plan = {}

    def real(v):
        return (v + root) % p

    top = 1
    while top < p:
        top <<= 1
    top >>= 1
    for v in range(p):
        start = top if v == 0 else ((v & -v) >> 1)
        sends = []
        m = start
        while m >= 1:
            if v + m < p:
                hi = min(v + 2 * m, p) - 1
                sends.append((real(v), real(v + m), v + m, hi))
            m >>= 1
        if sends:
            plan[real(v)] = sends
    return plan


def recv_num(plan, rank):
    return sum(1 for sends in plan.values() for (_, dst, _, _) in sends if dst == rank)
