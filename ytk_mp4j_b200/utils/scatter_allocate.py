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


class ScatterAllocate:
    """Name-compatible face of J/utils/ScatterAllocate.java (allocate :58-95, recvNum :114-131, main :36-56)."""
    allocate = staticmethod(allocate)
    recvNum = staticmethod(recv_num)

    @staticmethod
    def main(argv=None):
        """The reference's brute-force self-check: every root of a 13-rank job, every non-root rank
        receives exactly one message and the root none."""
        p = 13
        for root in range(p):
            plan = allocate(p, root)
            for rank in range(p):
                assert recv_num(plan, rank) == (0 if rank == root else 1), (root, rank)
        print("ScatterAllocate: %d roots ok" % p)
        return 0


def split(start, end, out):
    """Halving schedule of the rank range [start, end] (inclusive): the first rank of a range hands
    the upper half to that half's first rank, then both halves recurse. Appends
    (src, dst, rank_from, rank_to_inclusive) to `out` (behaviour of ScatterAllocate.split, :97-111)."""
    if start >= end:
        return out
    upper = start + (end - start + 1) // 2
    out.append((start, upper, upper, end))
    split(start, upper - 1, out)
    split(upper, end, out)
    return out


def allocate_halving(slave_num, root):
    """The reference's own scatter plan (ScatterAllocate.allocate, :58-95): the root first seeds rank 0
    with ranks [0, mid-1] and rank mid with [mid, p-1] (unless it is that rank itself), then both halves
    follow the halving schedule. For root not in {0, mid} the root's own slice travels out and comes
    back - reproduced here for the oracle, avoided by `allocate` above."""
    p, mid = slave_num, slave_num // 2
    sends = []
    if p > 1:
        if root != 0:
            sends.append((root, 0, 0, mid - 1))
        if root != mid:
            sends.append((root, mid, mid, p - 1))
        split(0, mid - 1, sends)
        split(mid, p - 1, sends)
    plan = {}
    for s in sends:
        plan.setdefault(s[0], []).append(s)
    return plan


ScatterAllocate.split = staticmethod(split)
ScatterAllocate.allocateHalving = staticmethod(allocate_halving)
