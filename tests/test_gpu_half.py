"""bf16 / fp16 dense collectives: every reduction step is fp32 math rounded to nearest-even, applied in
ascending rank order - so the result must equal, bit for bit, the same chain written in PyTorch."""
import pytest
import torch

from ytk_mp4j_b200 import _native as N

pytestmark = pytest.mark.gpu

TYPES = [(N.BF16, torch.bfloat16), (N.F16, torch.float16)]


@pytest.fixture(scope="module")
def planes():
    from ytk_mp4j_b200.parallel.gpu_plane import EmuPlane
    cache = {}

    def get(world):
        if world not in cache:
            cache[world] = EmuPlane(world, 0, heap_bytes=128 << 20, staging_bytes=4 << 20)
        return cache[world]
    yield get
    for p in cache.values():
        p.close()


def chain(parts, op):
    acc = parts[0].clone()
    for x in parts[1:]:
        a, b = acc.float(), x.float()
        r = {N.SUM: a + b, N.PROD: a * b, N.MAX: torch.maximum(a, b), N.MIN: torch.minimum(a, b)}[op]
        acc = r.to(parts[0].dtype)
    return acc


@pytest.mark.parametrize("code,tdt", TYPES)
@pytest.mark.parametrize("world", [2, 3, 8])
@pytest.mark.parametrize("op", [N.SUM, N.MAX, N.MIN, N.PROD])
def test_allreduce_matches_fp32_chain(planes, code, tdt, world, op):
    plane = planes(world)
    for n, f, symmetric in ((9, 0, False), (4099, 3, False), ((1 << 20) + 6, 2, True)):
        g = torch.Generator().manual_seed(n + world + op)
        host = [(torch.randn(n, generator=g) * (0.5 if op == N.PROD else 3.0)).to(tdt) for _ in range(world)]
        if symmetric:
            dev = plane.empty(n, code)
            for d, h in zip(dev, host):
                d.copy_(h)
        else:
            dev = [h.cuda() for h in host]
        plane.allreduce(dev, op, f, n)
        torch.cuda.synchronize()
        plane.check()
        want = chain([h[f:].cuda() for h in host], op)
        for r in (0, world - 1):
            assert torch.equal(dev[r][f:].view(torch.int16), want.view(torch.int16)), (n, r, plane.last())
            assert torch.equal(dev[r][:f].cpu().view(torch.int16), host[r][:f].view(torch.int16))


@pytest.mark.parametrize("code,tdt", TYPES)
def test_reduce_scatter_allgather_broadcast(planes, code, tdt):
    world, n = 4, 100003
    plane = planes(world)
    g = torch.Generator().manual_seed(7)
    host = [torch.randn(n, generator=g).to(tdt) for _ in range(world)]
    counts = [n // world] * (world - 1) + [n - (n // world) * (world - 1)]
    froms = [sum(counts[:q]) for q in range(world)]
    tos = [f + c for f, c in zip(froms, counts)]
    want = chain([h.cuda() for h in host], N.SUM)
    dev = [h.cuda() for h in host]
    plane.reduce_scatter(dev, N.SUM, 0, counts)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(dev[r][froms[r]:tos[r]].view(torch.int16), want[froms[r]:tos[r]].view(torch.int16)), r
    plane.allgather(dev, froms, tos)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(dev[r].view(torch.int16), want.view(torch.int16)), r
    dev = [h.cuda() for h in host]
    plane.broadcast(dev, 5, n, 2)
    torch.cuda.synchronize()
    assert torch.equal(dev[0][5:].view(torch.int16), host[2][5:].cuda().view(torch.int16))
    assert torch.equal(dev[0][:5].cpu().view(torch.int16), host[0][:5].view(torch.int16))
