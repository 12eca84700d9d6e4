"""Real multi-process GPU path: one process per GPU, symmetric heap mapped across processes (VMM fd
passing or CUDA IPC), signal pads over NVLink. Needs >= 2 GPUs; the single-GPU box skips it and
relies on the emulated-rank tests for the kernels plus the CPU tests for the control plane."""
import pytest
import torch

import workers_gpu
from ytk_mp4j_b200.parallel.local_cluster import run_cluster

pytestmark = pytest.mark.gpu


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("p", [2, 4, 8])
def test_process_collectives_on_gpus(p):
    if _ngpu() < p:
        pytest.skip("needs %d GPUs" % p)
    res, code, logs = run_cluster(p, workers_gpu.process_gpu, (), timeout=120)
    assert code == 0 and all(r["ok"] for r in res), res


@pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("p,T", [(2, 4), (4, 2)])
def test_thread_allreduce_on_gpus(p, T):
    if _ngpu() < p:
        pytest.skip("needs %d GPUs" % p)
    res, code, logs = run_cluster(p, workers_gpu.thread_gpu, (), threads=T, timeout=120)
    assert code == 0 and all(all(x) for x in res)


def test_single_process_comm_on_gpu():
    # p = 1: every collective short-circuits, thread collectives still run the intra-GPU kernel
    res, code, logs = run_cluster(1, workers_gpu.thread_gpu, (), threads=4, timeout=120)
    assert code == 0 and all(all(x) for x in res)


@pytest.mark.parametrize("p,T", [(1, 0), (1, 3), (2, 0), (2, 2), (4, 2)])
def test_check_matrix_on_cuda_tensors(p, T):
    """CommCheckTool's matrix with device=cuda (the reference's 128 check classes, arrays of the six primitive
    types) on 1, 2 or 4 GPUs, process level (T = 0) and thread level."""
    if _ngpu() < p:
        pytest.skip("needs %d GPUs" % p)
    res, code, logs = run_cluster(p, workers_gpu.check_matrix_cuda, (), threads=T, timeout=600)
    assert code == 0, logs[-20:]
    assert all(all(x) if isinstance(x, list) else x for x in res)


@pytest.mark.parametrize("p", [1, 2, 4])
def test_lr_and_gbdt_consumers_match_single_process_reference(p):
    """models/lr.py and models/gbdt.py end to end (the reference's named consumers, README.md:268-278): the loss
    curve of the data-parallel runs equals a single-process fp32 run on the concatenated data."""
    if _ngpu() < p:
        pytest.skip("needs %d GPUs" % p)
    res, code, logs = run_cluster(p, workers_gpu.models_worker, (), timeout=180)
    assert code == 0
    for out in res:
        assert out["dense_ok"], (out["dense_ours"], out["dense_ref"])
        assert out["sparse_ok"], (out["sparse_ours"], out["sparse_ref"])
        assert out["gbdt_ok"], out["gbdt"]


@pytest.mark.parametrize("T", [0, 3])
def test_cuda_tensors_without_a_gpu_plane_go_through_the_host_plane(T):
    """Ranks on several machines cannot share a GPU plane (one NVLink domain). MP4J_GPU_PLANE=0 puts a one-machine job
    in that state: collectives on CUDA tensors are staged through the host plane, init_gpu refuses with a reason."""
    fn = workers_gpu.staged_thread if T else workers_gpu.staged_process
    res, code, logs = run_cluster(2, fn, (), threads=T, timeout=120, env={"MP4J_GPU_PLANE": "0"})
    assert code == 0
    assert all(all(x) if isinstance(x, list) else x for x in res)
