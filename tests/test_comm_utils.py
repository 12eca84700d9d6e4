"""CommUtils / ScatterAllocate self-checks (the reference only has main() printouts:
J/utils/CommUtils.java:224-238, J/utils/ScatterAllocate.java:36-56)."""
import pytest

from ytk_mp4j_b200 import CommUtils, Mp4jException
from ytk_mp4j_b200.utils import scatter_allocate


def test_process_partition_matches_reference_rule():
    assert CommUtils.createProcessArrayFroms(101, 4) == [0, 25, 50, 75]
    assert CommUtils.createProcessArrayTos(101, 4) == [25, 50, 75, 101]


def test_thread_partition():
    f = CommUtils.createThreadArrayFroms(101, 4, 2)
    t = CommUtils.createThreadArrayTos(101, 4, 2)
    assert f == [[0, 12], [25, 37], [50, 62], [75, 88]]
    assert t == [[12, 25], [37, 50], [62, 75], [88, 101]]
    CommUtils.isfromsTosLegal(f, t, 2)
    assert CommUtils.getProcessFroms(f) == [0, 25, 50, 75]
    assert CommUtils.getProcessTos(t) == [25, 50, 75, 101]


def test_counts_helpers():
    assert CommUtils.getFromsFromCount(3, [2, 0, 5], 3) == [3, 5, 5]
    assert CommUtils.getTosFromCount(3, [2, 0, 5], 3) == [5, 5, 10]


@pytest.mark.parametrize("froms,tos", [([0, 5], [6, 9]), ([-1], [3]), ([4], [2]), ([0, 1], [1])])
def test_illegal_ranges(froms, tos):
    with pytest.raises(Mp4jException):
        CommUtils.isfromsTosLegal(froms, tos)


def test_illegal_counts_and_from_to():
    with pytest.raises(Mp4jException):
        CommUtils.isFromCountsLegal(-1, [1])
    with pytest.raises(Mp4jException):
        CommUtils.isFromCountsLegal(0, [[1, -1]])
    with pytest.raises(Mp4jException):
        CommUtils.isFromToLegal(5, 4)


@pytest.mark.parametrize("p", [1, 2, 3, 5, 8, 13])
def test_scatter_plan_every_rank_receives_once(p):
    # the reference brute-forces every root for p=13 and asserts recvNum <= 1
    for root in range(p):
        plan = scatter_allocate.allocate(p, root)
        for r in range(p):
            assert scatter_allocate.recv_num(plan, r) == (0 if r == root else 1)
        covered = sorted(v for sends in plan.values() for (_, _, lo, hi) in sends for v in range(lo, hi + 1)
                         if any(True for _ in [0]))
        # every relative rank 1..p-1 is the head of exactly one delivered range
        heads = sorted(lo for sends in plan.values() for (_, _, lo, _) in sends)
        assert heads == list(range(1, p))


def test_reference_halving_scatter_plan_delivers_every_slice():
    """ScatterAllocate.allocate of the reference (seed rank 0 and rank p/2, then halve): every rank ends
    up with its own slice, nobody but (possibly) the root receives twice - including the roots for which
    the root's own slice makes a round trip."""
    from ytk_mp4j_b200.utils.scatter_allocate import ScatterAllocate, allocate_halving, recv_num
    assert ScatterAllocate.main() == 0
    for p in (1, 2, 3, 5, 8, 13):
        for root in range(p):
            plan = allocate_halving(p, root)
            have = {root: set(range(p))}
            pending = [s for sends in plan.values() for s in sends]
            while pending:
                ready = [s for s in pending if s[0] in have and set(range(s[2], s[3] + 1)) <= have[s[0]]]
                assert ready, (p, root, pending)
                for s in ready:
                    have.setdefault(s[1], set()).update(range(s[2], s[3] + 1))
                    pending.remove(s)
            assert all(r in have.get(r, ()) for r in range(p)), (p, root)
            assert all(recv_num(plan, r) <= 1 for r in range(p) if r != root), (p, root)


def test_ranges_beyond_the_array_are_rejected():
    """A `to` / `tos` / `from + counts` beyond the array length must raise instead of letting native code
    read past the buffer (the reference throws ArrayIndexOutOfBoundsException)."""
    import pytest
    from ytk_mp4j_b200.exception import Mp4jException
    CommUtils.isFromToLegal(0, 10, 10)
    with pytest.raises(Mp4jException):
        CommUtils.isFromToLegal(0, 11, 10)
    CommUtils.isfromsTosLegal([0, 4], [4, 10], size=10)
    with pytest.raises(Mp4jException):
        CommUtils.isfromsTosLegal([0, 4], [4, 12], size=10)
    with pytest.raises(Mp4jException):
        CommUtils.isfromsTosLegal([[0, 2], [4, 6]], [[2, 4], [6, 11]], 2, size=10)
    CommUtils.isFromCountsLegal(2, [4, 4], 10)
    with pytest.raises(Mp4jException):
        CommUtils.isFromCountsLegal(3, [4, 4], 10)
    with pytest.raises(Mp4jException):
        CommUtils.isFromCountsLegal(0, [[4, 4], [2, 1]], 10)
