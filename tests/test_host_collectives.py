"""Multi-process CPU tests: control plane + C++ host collectives + generic (String/Object/map)
path + ThreadCommSlave, on local ranks over loopback TCP (the integration layer the reference
only tests on a real cluster)."""
import os

import pytest

import workers
from ytk_mp4j_b200.parallel.local_cluster import run_cluster


@pytest.mark.parametrize("p,dtype,n,root", [
    (1, "double", 100, 0),
    (2, "double", 1001, 1),
    (3, "float", 1000, 2),
    (4, "long", 103, 3),
    (4, "int", 4096, 0),
    (3, "short", 50, 1),
    (2, "byte", 77, 0),
    (4, "double", 3, 2),      # fewer elements than ranks: empty segments
])
@pytest.mark.parametrize("transport", ["shm", "tcp"])
def test_process_arrays(p, dtype, n, root, transport, monkeypatch):
    # shm: one gather-reduce-scatter pass over shared-memory windows (ranks share this machine);
    # tcp: the reference's ring / tree algorithms over sockets (the multi-host path)
    monkeypatch.setenv("MP4J_SHM", "1" if transport == "shm" else "0")
    res, code, _ = run_cluster(p, workers.process_arrays, (dtype, n, root, False, transport))
    assert res == [True] * p and code == 0


@pytest.mark.parametrize("transport", ["shm", "tcp"])
def test_process_arrays_large_pipelined(transport, monkeypatch):
    # tcp: > 1 MiB per segment exercises the chunked reduce-on-receive pipeline;
    # shm: a 256 KiB window forces many double-buffered chunks per collective
    monkeypatch.setenv("MP4J_SHM", "1" if transport == "shm" else "0")
    monkeypatch.setenv("MP4J_SHM_BYTES", str(256 << 10))
    res, code, _ = run_cluster(2, workers.process_arrays, ("double", 700_001, 1, False, transport))
    assert res == [True] * 2 and code == 0
    res, code, _ = run_cluster(3, workers.process_arrays, ("float", 300_007, 2, False, transport))
    assert res == [True] * 3 and code == 0


def test_loc_operators_end_to_end():
    res, code, _ = run_cluster(3, workers.process_loc_ops)
    assert res == [True] * 3 and code == 0


@pytest.mark.parametrize("p,root,compress", [(1, 0, False), (2, 1, False), (3, 2, True), (4, 0, False)])
def test_process_generic(p, root, compress):
    res, code, _ = run_cluster(p, workers.process_generic, (40, root, compress))
    assert res == [True] * p and code == 0


@pytest.mark.parametrize("p,T,dtype,root,rt", [(1, 1, "double", 0, 0), (1, 3, "float", 0, 2), (2, 2, "double", 1, 1),
                                               (3, 2, "int", 2, 0), (2, 4, "byte", 0, 3)])
def test_thread_arrays(p, T, dtype, root, rt):
    res, code, _ = run_cluster(p, workers.thread_arrays, (dtype, 120, root, rt), threads=T)
    assert all(all(x) for x in res) and code == 0


@pytest.mark.parametrize("p,T,root,rt", [(1, 2, 0, 1), (2, 2, 1, 0), (3, 2, 0, 1)])
def test_thread_maps(p, T, root, rt):
    res, code, _ = run_cluster(p, workers.thread_maps, (10, root, rt), threads=T)
    assert all(all(x) for x in res) and code == 0


def test_sorted_address_rank_rule():
    # no hints: rank = index of "host###port" in the sorted address list (SRV:188,199-206)
    res, code, _ = run_cluster(3, workers.control_plane, use_hints=False)
    assert sorted(res) == [0, 1, 2] and code == 0


def test_control_plane_logs_and_files(tmp_path):
    res, code, logs = run_cluster(3, workers.control_plane, master_kwargs={"work_dir": str(tmp_path)})
    assert res == [0, 1, 2] and code == 0
    text = "\n".join(logs)
    for r in range(3):
        assert "[rank=%d] hello from everyone" % r in text
    assert text.count("hello from rank0 only") == 1
    assert "[ERROR] [rank=1] an error line" in text
    assert (tmp_path / "mp4j_writefile_test.txt").read_text().strip() == "content-3"
    kills = [f for f in os.listdir(tmp_path) if f.startswith("kill_")]
    assert len(kills) == 1
    body = (tmp_path / kills[0]).read_text()
    assert body.count("kill -9") == 4  # three slaves + the master itself (SRV:331-342)


def test_failing_slave_ends_job_with_code_1():
    res, code, logs = run_cluster(2, workers.failing_slave)
    assert code == 1
    text = "\n".join(logs)
    assert "slave exception:" in text and "boom" in text


@pytest.mark.parametrize("native", ["1", "0"])
@pytest.mark.parametrize("p", [1, 2, 3, 4])
def test_string_key_maps_native_and_generic(p, native, monkeypatch):
    # native: C++ partition / ring / merge on key blobs; generic: the same algorithms in Python (pickle)
    monkeypatch.setenv("MP4J_NATIVE_MAPS", native)
    res, code, _ = run_cluster(p, workers.native_maps)
    assert res == [True] * p and code == 0


def test_gpu_plane_decision_is_the_same_on_every_rank():
    """The GPU plane is one NVLink domain; whether a job can have one is decided by all ranks together (hosts and
    MP4J_GPU_PLANE gathered through the master). One rank opting out must switch every rank to the host-staged path,
    otherwise the ranks would run different collectives."""
    res, code, _ = run_cluster(3, workers.plane_blocker, (-1,), timeout=60)
    assert code == 0 and all(r == "" for r in res), res
    res, code, _ = run_cluster(3, workers.plane_blocker, (1,), timeout=60)
    assert code == 0 and all(r == "MP4J_GPU_PLANE=0" for r in res), res
    res, code, _ = run_cluster(9, workers.plane_blocker, (-1,), timeout=120)  # more ranks than one NVLink domain holds
    assert code == 0 and all("at most 8" in r for r in res), res


@pytest.mark.parametrize("p,transport,window,n", [(2, "shm", 64 << 10, 40_003), (4, "shm", 64 << 10, 40_003),
                                                  (3, "shm", 64 << 20, 40_003), (3, "tcp", 0, 40_003),
                                                  (5, "shm", 128 << 10, 40_003),
                                                  (4, "tcp", 0, 1_501),    # <= 32 KiB: the direct all-pairs forms
                                                  (3, "tcp", 0, 4_099)])   # around the threshold: both forms mixed
def test_ragged_ranges_against_numpy(p, transport, window, n, monkeypatch):
    """Randomised ragged ranges (empty, unequal, with gaps) through every array collective. The small windows make
    the shared-memory plane take many passes, each over a part of every slice."""
    monkeypatch.setenv("MP4J_SHM", "1" if transport == "shm" else "0")
    if window:
        monkeypatch.setenv("MP4J_SHM_BYTES", str(window))
    res, code, _ = run_cluster(p, workers.ragged_arrays, (1234 + p, n), timeout=180)
    assert res == [True] * p and code == 0, res


@pytest.mark.parametrize("p,T,transport", [(1, 3, "shm"), (2, 2, "shm"), (3, 3, "tcp"), (2, 4, "shm"), (2, 3, "shm-solo")])
def test_ragged_thread_ranges_against_numpy(p, T, transport, monkeypatch):
    # shm: the T threads run the process step as a team; shm-solo (MP4J_THREAD_TEAM=0): the delegate alone; tcp: the
    # delegate runs the ring / tree, the other threads fall through to the barrier
    monkeypatch.setenv("MP4J_SHM", "0" if transport == "tcp" else "1")
    monkeypatch.setenv("MP4J_THREAD_TEAM", "0" if transport == "shm-solo" else "1")
    monkeypatch.setenv("MP4J_SHM_BYTES", str(64 << 10))
    res, code, _ = run_cluster(p, workers.ragged_thread_arrays, (77 + p * T, 20_011), threads=T, timeout=180)
    assert code == 0 and all(all(x is True for x in row) for row in res), res


@pytest.mark.parametrize("p,transport", [(20, "shm"), (36, "tcp")])
def test_many_slaves(p, transport, monkeypatch):
    """More ranks than cores; 36 is beyond the shared-memory plane's 32-rank masks, so every rank must fall back to
    the TCP ring together."""
    monkeypatch.setenv("MP4J_SHM", "1")
    res, code, _ = run_cluster(p, workers.many_slaves, (), timeout=240)
    assert code == 0 and res == [True] * p


@pytest.mark.parametrize("ids,n", [("0,0,1,1", 100_003), ("0,0,0,1,1,1", 20_011), ("a,b,a,b,a,b,a,b", 300_007),
                                   ("0,0,1,1,2,2", 5_000)])
def test_hierarchical_allreduce_across_virtual_machines(ids, n, monkeypatch):
    """No job-wide shared memory (the ranks sit on different 'machines'), but one plane per machine: allreduce =
    shared-memory reduce-scatter + TCP between the machines + shared-memory allgather. Interleaved rank placement
    (a,b,a,b,...) checks that groups come from the machine ids, not from rank ranges."""
    p = len(ids.split(","))
    monkeypatch.setenv("MP4J_SHM_BYTES", str(256 << 10))  # several passes per plane
    res, code, _ = run_cluster(p, workers.hierarchical, (n,), timeout=180, env={"MP4J_NODE_IDS": ids})
    assert code == 0 and res == [True] * p, res


def test_uneven_machines_fall_back_to_the_flat_tcp_paths():
    res, code, _ = run_cluster(5, workers.plane_blocker, (-1,), timeout=60, env={"MP4J_NODE_IDS": "0,0,0,1,1"})
    assert code == 0  # 3 + 2 ranks: no hierarchy, and nothing hangs while deciding so
    assert all("2 machines" in r for r in res), res  # and no GPU plane across "machines" either


@pytest.mark.parametrize("ids", ["0,0,1,1", "a,b,a,b,a,b", "0,0,0,1,1,1,2,2,2"])
def test_ragged_ranges_on_virtual_machines(ids, monkeypatch):
    """Ragged ranges through every array collective when allreduce, reduce-scatter and allgather are hierarchical
    (per-machine shared memory + TCP rails) and the rooted ones run on the flat mesh."""
    p = len(ids.split(","))
    monkeypatch.setenv("MP4J_SHM_BYTES", str(64 << 10))
    res, code, _ = run_cluster(p, workers.ragged_arrays, (4321 + p, 40_003), timeout=240, env={"MP4J_NODE_IDS": ids})
    assert code == 0 and res == [True] * p, res
