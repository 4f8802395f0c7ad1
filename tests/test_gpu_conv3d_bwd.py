"""GPU parity of the conv backward kernels against torch-CPU autograd on the same bf16-rounded operands."""
import pytest
import torch
import torch.nn.functional as F

from oracle import model_torch as OM

pytestmark = pytest.mark.gpu


def _bf16_exact(shape, gen, scale=1.0):
    return (torch.randn(shape, generator=gen) * scale).to(torch.bfloat16).float()


@pytest.mark.parametrize("idx,Din,H,W", [(1, 5, 24, 16), (2, 3, 40, 24), (0, 10, 17, 16)])
def test_wgrad_matches_autograd(cuda_device, idx, Din, H, W):
    """dW of the three layer geometries (model.py:207-209); H = 17 exercises a partial tile row."""
    from voxelnet_pytorch_b200 import ops

    cin, _, stride, padding = OM.MIDDLE_LAYERS[idx]
    g = torch.Generator().manual_seed(40 + idx)
    B = 2
    x = _bf16_exact((B, Din, H, W, cin), g)
    Dout = (Din + 2 * padding[0] - 3) // stride[0] + 1
    dy = _bf16_exact((B, Dout, H, W, 64), g, 0.5)
    w = torch.zeros((64, cin, 3, 3, 3), requires_grad=True)
    y = F.conv3d(x.permute(0, 4, 1, 2, 3), w, None, stride=stride, padding=padding)
    y.backward(dy.permute(0, 4, 1, 2, 3))
    want = w.grad
    got = ops.conv3d_wgrad(x.to(cuda_device).to(torch.bfloat16), dy.to(cuda_device).to(torch.bfloat16), stride[0], padding[0]).cpu()
    err = float((got - want).abs().max() / want.abs().max())
    assert got.shape == want.shape and err < 2e-4, err


@pytest.mark.parametrize("idx,Din,H,W", [(1, 5, 24, 16), (2, 3, 40, 24), (2, 3, 17, 8)])
def test_dgrad_matches_autograd(cuda_device, idx, Din, H, W):
    """dX of layers 2 and 3 (64 input channels): the forward kernel on the transposed depth geometry."""
    from voxelnet_pytorch_b200 import ops

    cin, _, stride, padding = OM.MIDDLE_LAYERS[idx]
    g = torch.Generator().manual_seed(50 + idx)
    B = 2
    Dout = (Din + 2 * padding[0] - 3) // stride[0] + 1
    w = _bf16_exact((64, cin, 3, 3, 3), g, 0.05)
    dy = _bf16_exact((B, Dout, H, W, 64), g)
    x = torch.zeros((B, cin, Din, H, W), requires_grad=True)
    y = F.conv3d(x, w, None, stride=stride, padding=padding)
    y.backward(dy.permute(0, 4, 1, 2, 3))
    want = x.grad.permute(0, 2, 3, 4, 1)
    packed = ops.conv3d_pack_weights_dgrad(w.to(cuda_device))
    got = ops.conv3d_dgrad(dy.to(cuda_device).to(torch.bfloat16), packed, Din, stride[0], padding[0]).float().cpu()
    err = float((got - want).abs().max() / want.abs().max())
    assert got.shape == want.shape and err < 6e-3, err  # bf16 storage of dx


def test_wgrad_skips_dead_boxes_exactly(cuda_device):
    """First layer on a sparse voxel grid: with the live-box table of vn_conv3d_mark_boxes the units whose
    input boxes hold only zeros are skipped; the result equals the unflagged run up to summation order."""
    from voxelnet_pytorch_b200 import ops

    g = torch.Generator().manual_seed(70)
    B, D, H, W = 2, 10, 40, 32
    occ = torch.rand((B, D, H, W), generator=g) < 0.002
    occ[0, 0, 0, 0] = occ[1, 9, H - 1, W - 1] = True
    coords = occ.nonzero().contiguous().to(cuda_device)
    x = torch.zeros((B, D, H, W, 128))
    x[occ] = _bf16_exact((int(occ.sum()), 128), g)
    dy = _bf16_exact((B, 5, H, W, 64), g, 0.5)
    xd, dyd = x.to(cuda_device).to(torch.bfloat16), dy.to(cuda_device).to(torch.bfloat16)
    flags = ops.conv3d_new_box_flags(B, D, H, W, cuda_device)
    ops.conv3d_mark_boxes(coords, B, D, H, W, [(2, 1)], [flags])
    assert 0 < int(flags.sum()) < flags.numel()
    dense = ops.conv3d_wgrad(xd, dyd, 2, 1)
    sparse = ops.conv3d_wgrad(xd, dyd, 2, 1, box_flags=flags)
    err = float((dense - sparse).abs().max() / dense.abs().max())
    assert err < 1e-5, err
