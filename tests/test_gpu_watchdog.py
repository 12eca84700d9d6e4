"""Device-side watchdog: a rank that never arrives must not hang the others, the kernel must
store nothing after the failed wait, and the host must hear about it - at check() and at the next
collective - instead of returning garbage with rc 0."""
import pytest
import torch

from ytk_mp4j_b200 import Mp4jException
from ytk_mp4j_b200 import _native as N

pytestmark = pytest.mark.gpu


def test_missing_peer_raises_and_poisons_the_plane():
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    plane = EmuPlane(2, 0, heap_bytes=192 << 20, staging_bytes=4 << 20)
    try:
        n = 1 << 20
        ts = [torch.full((n,), float(r + 1), device="cuda") for r in range(2)]
        plane.allreduce(ts, N.SUM, 0, n, algo=N.ALGO_TWOSHOT)  # healthy call first
        torch.cuda.synchronize()
        plane.check()
        assert torch.all(ts[0] == 3.0)
        before = ts[0].clone()
        plane.set_option("timeout_ms", 200)
        plane.set_option("dead_rank", 1)  # rank 1's blocks return at once: it "died"
        plane.allreduce(ts, N.SUM, 0, n, algo=N.ALGO_TWOSHOT)
        torch.cuda.synchronize()  # returns: the waits are bounded
        assert torch.equal(ts[0], before), "a kernel whose peer is missing must not store"
        with pytest.raises(Mp4jException, match="did not arrive"):
            plane.check()
        plane.set_option("dead_rank", -1)
        with pytest.raises(Mp4jException, match="unusable"):
            plane.allreduce(ts, N.SUM, 0, n, algo=N.ALGO_TWOSHOT)
    finally:
        plane.close()
