"""GPU sparse-map collectives (allreduceMap / reduceScatterMap / allgatherMap on tensors) against
a dictionary-based PyTorch/CPU reference; value patterns of the reference's map checks
(J/check/checkdouble/ProcessAllReduceCheck.java:86-116: shared keys + one private key per rank)."""
import pytest
import torch

from ytk_mp4j_b200 import _native as N

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def planes():
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    cache = {}

    def get(world):
        if world not in cache:
            cache[world] = EmuPlane(world, 0, heap_bytes=256 << 20, staging_bytes=4 << 20)
        return cache[world]
    yield get
    for p in cache.values():
        p.close()


def ref_reduce(keys, vals, op):
    out = {}
    for k, v in zip(keys, vals):
        for kk, vv in zip(k.tolist(), v.tolist()):
            if kk in out:
                out[kk] = out[kk] + vv if op == N.SUM else (max(out[kk], vv) if op == N.MAX else min(out[kk], vv))
            else:
                out[kk] = vv
    return out


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_allreduce_map_reference_pattern(planes, world):
    plane = planes(world)
    obj = 5000
    keys, vals = [], []
    for r in range(world):
        k = torch.cat([torch.arange(obj, dtype=torch.int64), torch.tensor([-(r + 1)], dtype=torch.int64)])
        keys.append(k.cuda())
        vals.append(torch.ones(obj + 1, dtype=torch.float64).cuda())
    res = plane.map_collective("allreduce", keys, vals, N.SUM)
    plane.check()
    for r in range(world):
        k, v, _ = res[r]
        assert k.numel() == obj + world
        d = dict(zip(k.tolist(), v.tolist()))
        assert all(d[i] == world for i in range(obj))
        assert all(d[-(q + 1)] == 1 for q in range(world))


@pytest.mark.parametrize("dtype,op", [(torch.float32, N.SUM), (torch.int64, N.SUM), (torch.int32, N.MAX),
                                      (torch.float64, N.MIN), (torch.float32, N.MAX)])
def test_allreduce_map_random_keys_all_types(planes, dtype, op):
    world = 4
    plane = planes(world)
    g = torch.Generator().manual_seed(17)
    keys, vals = [], []
    for r in range(world):
        k = torch.randperm(40000, generator=g)[:12000].to(torch.int64) * 7919 - 1000  # unique per rank, overlapping
        v = torch.randint(-50, 50, (12000,), generator=g).to(dtype)
        keys.append(k)
        vals.append(v)
    want = ref_reduce(keys, vals, op)
    for rep in range(2):  # second call reuses buffers and the alternate count slots
        res = plane.map_collective("allreduce", [k.cuda() for k in keys], [v.cuda() for v in vals], op)
        for r in range(world):
            k, v, _ = res[r]
            got = dict(zip(k.tolist(), v.tolist()))
            assert len(got) == len(want) == k.numel()
            assert got == want
    plane.check()


def test_reduce_scatter_map_ownership_and_java_partition(planes):
    world = 4
    plane = planes(world)
    from ytk_mp4j_b200.comm.process import java_string_hash, key_partition
    names = ["feature_%d" % i for i in range(3000)]
    hashed = torch.tensor([java_string_hash(s) & 0xFFFFFFFF for s in names], dtype=torch.int64)
    keys = [hashed.cuda() for _ in range(world)]
    vals = [torch.full((3000,), float(r + 1), dtype=torch.float32).cuda() for r in range(world)]
    res = plane.map_collective("reduce_scatter", keys, vals, N.SUM, partition="java32")
    uniq = {}
    for s, h in zip(names, hashed.tolist()):
        uniq.setdefault(h, []).append(s)
    seen = 0
    for r in range(world):
        k, v, own = res[r]
        assert own == k.numel()
        seen += k.numel()
        for kk, vv in zip(k.tolist(), v.tolist()):
            # placement matches the reference's String.hashCode() % slaveNum (PCS:2066-2070)
            assert key_partition(uniq[kk][0], world) == r
            assert vv == len(uniq[kk]) * world * (world + 1) / 2
    assert seen == len(uniq)
    plane.check()


def test_allgather_map_rank_order(planes):
    world = 3 if False else 4
    plane = planes(world)
    obj = 1000
    keys = [torch.arange(r * obj, (r + 1) * obj, dtype=torch.int64).cuda() for r in range(world)]
    vals = [torch.arange(r * obj, (r + 1) * obj, dtype=torch.float64).cuda() for r in range(world)]
    res = plane.map_collective("allgather", keys, vals)
    for r in range(world):
        k, v, _ = res[r]
        assert torch.equal(k.cpu(), torch.arange(world * obj, dtype=torch.int64))
        assert torch.equal(v.cpu(), torch.arange(world * obj, dtype=torch.float64))
    plane.check()


def test_map_output_overflow_is_reported(planes):
    from ytk_mp4j_b200 import Mp4jException
    plane = planes(2)
    keys = [torch.arange(100, dtype=torch.int64).cuda() + 1000 * r for r in range(2)]
    vals = [torch.ones(100, dtype=torch.float32).cuda() for _ in range(2)]
    with pytest.raises(Mp4jException):
        plane.map_collective("allreduce", keys, vals, N.SUM, out_cap=150)


@pytest.mark.parametrize("dtype,op", [(torch.float32, N.SUM), (torch.int64, N.MAX), (torch.float64, N.MIN)])
@pytest.mark.parametrize("n", [1, 1000, 200000])
def test_map_reduce_local(planes, dtype, op, n):
    """Thread-level map merge (DOP:458-477) on device arrays: reduce-by-key with a scratch hash table."""
    plane = planes(1)
    g = torch.Generator().manual_seed(n)
    keys = torch.randint(-n, n, (n,), generator=g, dtype=torch.int64)
    vals = torch.randint(-9, 9, (n,), generator=g).to(dtype)
    want = ref_reduce([keys], [vals], op)
    for rep in range(2):
        k, v = plane.map_reduce_local(keys.cuda(), vals.cuda(), op)
        got = dict(zip(k.tolist(), v.tolist()))
        assert got == want and k.numel() == len(want)
