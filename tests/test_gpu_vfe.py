"""GPU parity: fused SVFE and per-layer VFE kernels vs the CPU fp32 oracle (rel 1e-4)."""
import numpy as np
import pytest
import torch

from oracle import model_torch as OM
from oracle import voxelize_np as OV
from tests.helpers import load_state, synth_frames

pytestmark = pytest.mark.gpu

RTOL = 1e-4  # north_star: "VFE outputs must be within 1e-4 relative in fp32"


def _close(got, want, what):
    got, want = got.double().cpu(), want.double()
    err = (got - want).abs()
    tol = RTOL * want.abs() + RTOL * float(want.abs().max()) * 1e-2
    assert bool((err <= tol).all()), "%s: max err %.3e (max |ref| %.3e)" % (what, float(err.max()), float(want.abs().max()))


@pytest.mark.parametrize("target", ["Car", "Pedestrian"])
def test_svfe_fused_matches_oracle(cuda_device, target):
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(1)
    feat = torch.from_numpy(OV.pcl_to_voxels_fast(synth_frames(1, 2)[0], target)["feature_buffer"])
    want = OM.svfe_voxelwise(feat, sd)
    net = vn.FeatureLearningNet(target).eval()
    load_state(net, sd, "")
    net = net.to(cuda_device)
    got = net.voxelwise(feat.to(cuda_device))
    _close(got, want, "svfe voxelwise")


def test_svfe_handles_arbitrary_rows(cuda_device):
    """No assumption about pad rows: random features incl. all-zero rows, rows whose max is 0
    (mask off although non-zero), fully populated voxels and voxels made of one repeated row."""
    import voxelnet_pytorch_b200 as vn

    g = torch.Generator().manual_seed(0)
    K, T = 300, 35
    feat = torch.randn((K, T, 7), generator=g)
    feat[:40, 5:] = 0.0                       # zero tails
    feat[40:80] = -feat[40:80].abs()          # all negative -> max < 0 -> mask on
    feat[80:120, :, :] = torch.minimum(feat[80:120], torch.zeros(()))  # max may be exactly 0 -> mask off
    feat[80:120, ::2, 3] = 0.0
    feat[120:160] = feat[120:160, :1].expand(-1, T, -1)  # one repeated row
    feat[160:170] = 0.0                       # completely empty voxel
    sd = OM.random_state_dict(2)
    want = OM.svfe_voxelwise(feat, sd)
    net = vn.FeatureLearningNet().eval()
    load_state(net, sd, "")
    got = net.to(cuda_device).voxelwise(feat.to(cuda_device))
    _close(got, want, "svfe voxelwise (arbitrary rows)")


@pytest.mark.parametrize("layer,cin,cout", [("vfe_1.", 7, 32), ("vfe_2.", 32, 128)])
def test_vfe_layer_dropin(cuda_device, layer, cin, cout):
    import voxelnet_pytorch_b200 as vn

    g = torch.Generator().manual_seed(3)
    K, T = 257, 35
    x = torch.randn((K, T, cin), generator=g)
    x[:, 20:] *= (torch.rand((K, 15, 1), generator=g) > 0.5)
    mask = torch.rand((K, T, 1), generator=g) > 0.3
    sd = OM.random_state_dict(4)
    want = OM.vfe_layer(x, mask, sd, layer)
    m = vn.VFELayer(cin, cout).eval()
    load_state(m, sd, layer)
    got = m.to(cuda_device)(x.to(cuda_device), mask.to(cuda_device))
    assert got.shape == (K, T, cout)
    _close(got, want, "vfe layer " + layer)


def test_feature_mask(cuda_device):
    from voxelnet_pytorch_b200 import ops

    feat = torch.from_numpy(OV.pcl_to_voxels_fast(synth_frames(1, 4)[0], "Car")["feature_buffer"])
    want = OM.feature_mask(feat)
    got = ops.feature_mask(feat.to(cuda_device))
    assert got.dtype == torch.bool and torch.equal(got.cpu(), want)


@pytest.mark.parametrize("cin,cout,T", [(7, 32, 35), (32, 128, 45)])
def test_single_layer_train_mode_matches_oracle(cuda_device, cin, cout, T):
    """VFELayer.forward alone under .train() (model.py:74-82): BatchNorm1d with the statistics of its own K*T input
    rows (pads included), running statistics updated like nn.BatchNorm1d; gradients of the stand-alone layer against
    autograd over the oracle."""
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(61)
    prefix = "vfe_1." if cin == 7 else "vfe_2."
    g = torch.Generator().manual_seed(62)
    x = torch.randn((97, T, cin), generator=g)
    mask = torch.rand((97, T, 1), generator=g) > 0.4
    layer = vn.VFELayer(cin, cout)
    load_state(layer, sd, prefix)
    layer = layer.to(cuda_device).train()
    new_stats = {}
    want = OM.vfe_layer(x, mask, sd, prefix, training=True, new_stats=new_stats)
    with torch.no_grad():
        got = layer(x.to(cuda_device), mask.to(cuda_device))
    assert got.shape == want.shape
    assert torch.allclose(got.cpu(), want, rtol=1e-4, atol=2e-5)
    assert torch.allclose(layer.bn.running_mean.cpu(), new_stats[prefix + "bn.running_mean"], rtol=1e-5, atol=1e-6)
    assert torch.allclose(layer.bn.running_var.cpu(), new_stats[prefix + "bn.running_var"], rtol=1e-5, atol=1e-6)
    assert int(layer.bn.num_batches_tracked) == 1
    # with gradients: the autograd node of the stand-alone layer
    leaves = {prefix + n: sd[prefix + n].clone().requires_grad_(True) for n in ("fcn.0.weight", "fcn.0.bias", "bn.weight", "bn.bias")}
    xo = x.clone().requires_grad_(True)
    up = torch.randn(want.shape, generator=g)
    OM.vfe_layer(xo, mask, {**sd, **leaves}, prefix, training=True).backward(up)
    xg = x.to(cuda_device).requires_grad_(True)
    layer.zero_grad()
    layer(xg, mask.to(cuda_device)).backward(up.to(cuda_device))
    assert torch.allclose(xg.grad.cpu(), xo.grad, rtol=1e-3, atol=1e-5)
    for name, p in (("fcn.0.weight", layer.fcn[0].weight), ("fcn.0.bias", layer.fcn[0].bias), ("bn.weight", layer.bn.weight), ("bn.bias", layer.bn.bias)):
        assert torch.allclose(p.grad.cpu(), leaves[prefix + name].grad, rtol=1e-3, atol=1e-4), name
