"""Fail-fast behaviour: a slave that vanishes in the middle of a job must make its peers'
next collective fail within seconds - on both host transports - instead of hanging them."""
import multiprocessing as mp

import pytest

import workers
from ytk_mp4j_b200 import CommMaster
from ytk_mp4j_b200.parallel.local_cluster import _worker


@pytest.mark.parametrize("transport,n", [("shm", 1000), ("tcp", 1000),   # tcp, 8 KB: the direct all-pairs exchange
                                         ("tcp", 2_000_000)])            # tcp, 16 MB: ring steps with a helper thread
def test_peer_death_is_detected(transport, n, tmp_path):
    env = {"MP4J_SHM": "1" if transport == "shm" else "0", "MP4J_PEER_TIMEOUT_S": "20"}
    ctx = mp.get_context("spawn")
    master = CommMaster(2, 0, quiet=True, heartbeat_timeout_s=30, work_dir=str(tmp_path)).start()
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(workers.dying_peer, i, 2, master.port, 0, (n,), q, env)) for i in range(2)]
    for p in procs:
        p.start()
    try:
        hint, rank, status, out = q.get(timeout=60)  # only rank 0 reports; rank 1 is gone
        assert rank == 0
        text = str(out)
        assert "failed-fast" in text or status != "ok", text  # an Mp4jException, not a hang and not success
        assert "no error" not in text
    finally:
        for p in procs:
            p.join(timeout=5)
            if p.is_alive():
                p.kill()
        master.shutdown()
        master.close()
    assert procs[1].exitcode == 7
