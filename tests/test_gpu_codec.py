"""K8: wire-compressed allreduce (bf16 / block-scaled fp8). Lossy by design, so: tolerance against
the lossless fp32 sum, bit-identical results on all ranks, and exactness where the format is exact."""
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


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("wire,rel", [(N.WIRE_BF16, 2 ** -7), (N.WIRE_FP8, 2 ** -2.5)])
@pytest.mark.parametrize("n,f", [(128, 0), (100003, 4), (1 << 20, 0)])
def test_wire_allreduce_tolerance_and_agreement(planes, world, wire, rel, n, f):
    plane = planes(world)
    g = torch.Generator().manual_seed(n + world)
    host = [torch.randn(n, generator=g) for _ in range(world)]
    dev = [h.cuda() for h in host]
    plane.allreduce_wire(dev, f, n, wire)
    torch.cuda.synchronize()
    plane.check()
    want = torch.stack([h[f:] for h in host]).sum(0)
    got0 = dev[0][f:].cpu()
    # per quantisation block the error is bounded by the block's amax times the format's step
    scale = torch.stack([h[f:].abs() for h in host]).sum(0).max().item()
    assert (got0 - want).abs().max().item() <= rel * scale * 2
    for r in range(world):
        assert torch.equal(dev[r][f:].cpu(), got0), "all ranks must decode identical bytes"
        assert torch.equal(dev[r][:f].cpu(), host[r][:f])


def test_wire_formats_are_exact_on_representable_values(planes):
    plane = planes(4)
    n = 4096
    for wire in (N.WIRE_BF16, N.WIRE_FP8):
        dev = [torch.full((n,), float(r + 1)).cuda() for r in range(4)]  # small integers: exact in both formats
        plane.allreduce_wire(dev, 0, n, wire)
        torch.cuda.synchronize()
        assert all(bool(torch.all(d == 10.0)) for d in dev), plane.last()
