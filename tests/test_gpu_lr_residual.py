"""Fused softmax-residual kernel of the LR consumer (csrc/gpu/lr_residual.cu) against plain PyTorch fp32."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,C", [(1, 2), (33, 7), (1000, 128), (4099, 1000), (65, 3072)])
def test_softmax_residual_transposed(rows, C):
    from ytk_mp4j_b200.parallel.gpu_plane import lr_softmax_residual
    g = torch.Generator().manual_seed(rows + C)
    logits = (torch.randn(rows, C, generator=g) * 3).to(torch.bfloat16)
    y = torch.randint(0, C, (rows,), generator=g)
    y[::7] = -1  # padding rows
    rt, loss = lr_softmax_residual(logits.cuda(), y.cuda())
    torch.cuda.synchronize()
    x = logits.float()
    logp = torch.log_softmax(x, 1)
    valid = y >= 0
    want = logp.exp()
    want[torch.arange(rows)[valid], y[valid]] -= 1.0
    want[~valid] = 0.0
    want_loss = float(-(logp[torch.arange(rows)[valid], y[valid]]).sum())
    got = rt.float().cpu().t()
    assert got.shape == (rows, C)
    assert (got - want).abs().max().item() <= 8e-3  # bf16 output: 2^-8 relative on values <= 1
    assert abs(float(loss) - want_loss) <= 1e-3 * abs(want_loss) + 1e-3
    # the loss accumulates: a second call adds to it
    lr_softmax_residual(logits.cuda(), y.cuda(), rt, loss)
    torch.cuda.synchronize()
    assert abs(float(loss) - 2 * want_loss) <= 2e-3 * abs(want_loss) + 2e-3
