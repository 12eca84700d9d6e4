"""K12: tcgen05 GEMM (X^T R) fused with the reduce-scatter, against an fp32 PyTorch matmul of the
same bf16 inputs + a reference reduce-scatter (sum over ranks, row ranges per owner)."""
import pytest
import torch

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


@pytest.mark.parametrize("world,F,C,K", [
    (1, 128, 64, 64),          # one tile, one k-block
    (1, 256, 128, 512),        # pipeline wraps (8 k-blocks, 4 stages)
    (1, 1000, 32, 4104),       # ragged F (OOB rows), K not a multiple of 64, split-K
    (2, 512, 128, 2048),
    (4, 1024, 256, 1024),
    (8, 4096, 16, 8192),
])
@pytest.mark.parametrize("split", ["atomic", "cluster", "atomic-2mt"])
def test_lr_xtr_reduce_scatter(planes, world, F, C, K, split, monkeypatch):
    # read by the launcher at every call; "2mt" = two 128-row tiles per CTA sharing each staged B tile
    monkeypatch.setenv("MP4J_LR_SPLIT", split.split("-")[0])
    monkeypatch.setenv("MP4J_LR_MT", "2" if split.endswith("2mt") else "1")
    plane = planes(world)
    g = torch.Generator().manual_seed(F + C + K)
    xts = [(torch.randn(F, K, generator=g) * 0.5).to(torch.bfloat16) for _ in range(world)]
    rts = [(torch.randn(C, K, generator=g) * 0.5).to(torch.bfloat16) for _ in range(world)]
    for rep in range(2):
        shards, fpr = plane.lr_xtr_reduce_scatter([x.cuda() for x in xts], [r.cuda() for r in rts])
        torch.cuda.synchronize()
        plane.check()
        want = sum(x.float() @ r.float().t() for x, r in zip(xts, rts))  # [F, C] fp32 reference
        assert fpr == (F + world - 1) // world
        for q in range(world):
            lo, hi = q * fpr, min(F, (q + 1) * fpr)
            got = shards[q][: hi - lo].cpu()
            err = (got - want[lo:hi]).abs().max().item()
            scale = want.abs().max().item()
            assert err <= 2e-3 * scale + 1e-3, (rep, q, err, scale)
