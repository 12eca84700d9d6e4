"""K13: fused GBDT histogram build + reduce-scatter against a PyTorch index_add_ reference."""
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


def ref_hist(bins_t, grad, hess):
    F, rows = bins_t.shape
    out = torch.zeros(F, 256, 2, dtype=torch.float64)
    idx = bins_t.long()
    for f in range(F):
        out[f, :, 0].index_add_(0, idx[f], grad.double())
        out[f, :, 1].index_add_(0, idx[f], hess.double())
    return out


@pytest.mark.parametrize("world,F,rows", [(1, 8, 4096), (1, 13, 10007), (2, 64, 50000), (4, 30, 33333), (8, 100, 20000)])
def test_gbdt_hist_reduce_scatter(planes, world, F, rows):
    plane = planes(world)
    g = torch.Generator().manual_seed(F * rows)
    pitch = (rows + 15) // 16 * 16
    bins, grads, hesss, want = [], [], [], torch.zeros(F, 256, 2, dtype=torch.float64)
    for r in range(world):
        b = torch.randint(0, 256, (F, rows), generator=g, dtype=torch.uint8)
        gr = torch.randint(-8, 8, (rows,), generator=g).float()   # integer-valued: fp32 sums are exact
        he = torch.randint(0, 4, (rows,), generator=g).float()
        want += ref_hist(b, gr, he)
        buf = torch.zeros(F, pitch, dtype=torch.uint8)
        buf[:, :rows] = b
        bins.append(buf.cuda()[:, :rows])
        grads.append(gr.cuda())
        hesss.append(he.cuda())
    for rep in range(2):
        shards, fpr = plane.gbdt_hist_reduce_scatter(bins, grads, hesss)
        torch.cuda.synchronize()
        plane.check()
        assert fpr == (F + world - 1) // world
        for q in range(world):
            lo, hi = q * fpr, min(F, (q + 1) * fpr)
            if hi > lo:
                assert torch.equal(shards[q][: hi - lo].cpu().double(), want[lo:hi]), (rep, q)
