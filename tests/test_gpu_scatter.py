"""GPU parity: sparse->dense scatter and FeatureLearningNet.forward vs the oracle."""
import numpy as np
import pytest
import torch

from oracle import model_torch as OM
from tests.helpers import load_state, oracle_batch, synth_frames

pytestmark = pytest.mark.gpu


def test_scatter_f32_exact_and_sums_duplicates(cuda_device):
    from voxelnet_pytorch_b200 import ops

    g = torch.Generator().manual_seed(0)
    B, D, H, W, C = 2, 10, 200, 240, 128
    K = 500
    coords = torch.stack([torch.randint(0, B, (K,), generator=g), torch.randint(0, D, (K,), generator=g),
                          torch.randint(0, H, (K,), generator=g), torch.randint(0, W, (K,), generator=g)], 1)
    coords[10] = coords[11]  # duplicate coordinate: COO to_dense sums
    vw = torch.randn((K, C), generator=g)
    want = OM.scatter_dense(vw, coords, B, "Pedestrian")
    dense = torch.empty((B, D, H, W, C), dtype=torch.float32, device=cuda_device)
    ops.fill_zero(dense)
    ops.scatter_dense_f32(vw.to(cuda_device), coords.to(cuda_device), dense)
    got = dense.cpu()
    dup = tuple(int(v) for v in coords[10])
    assert torch.allclose(got[dup], want[dup], rtol=1e-6, atol=1e-6)
    got[dup] = 0
    want[dup] = 0
    assert torch.equal(got, want)


def test_scatter_bf16_write_then_clear(cuda_device):
    from voxelnet_pytorch_b200 import ops

    g = torch.Generator().manual_seed(1)
    B, D, H, W, C = 1, 10, 64, 48, 128
    K = 300
    flat = torch.randperm(B * D * H * W, generator=g)[:K]
    coords = torch.stack([flat // (D * H * W), (flat // (H * W)) % D, (flat // W) % H, flat % W], 1)
    vw = torch.randn((K, C), generator=g)
    dense = torch.empty((B, D, H, W, C), dtype=torch.bfloat16, device=cuda_device)
    ops.fill_zero(dense)
    live = torch.tensor([K - 7], dtype=torch.int32, device=cuda_device)  # device-side live count
    ops.scatter_dense_bf16(vw.to(cuda_device), coords.to(cuda_device), dense, n_voxels_dev=live)
    want = torch.zeros((B, D, H, W, C), dtype=torch.bfloat16)
    c = coords[: K - 7]
    want[c[:, 0], c[:, 1], c[:, 2], c[:, 3]] = vw[: K - 7].to(torch.bfloat16)
    assert torch.equal(dense.cpu(), want)
    ops.scatter_dense_bf16(None, coords.to(cuda_device), dense, clear=True, n_voxels_dev=live)
    assert int(dense.view(torch.int16).ne(0).sum()) == 0


def test_feature_learning_net_forward(cuda_device):
    import voxelnet_pytorch_b200 as vn

    frames = synth_frames(2, 20)
    frames = [f[:30000] for f in frames]
    _, feats, _, coords = oracle_batch(frames, "Pedestrian")
    sd = OM.random_state_dict(5)
    tf = [torch.from_numpy(f) for f in feats]
    tc = [torch.from_numpy(c) for c in coords]
    want = OM.feature_learning_net(tf, tc, sd, "Pedestrian")
    net = vn.FeatureLearningNet("Pedestrian").eval()
    load_state(net, sd, "")
    net = net.to(cuda_device)
    got = net([t.to(cuda_device) for t in tf], [t.to(cuda_device) for t in tc]).cpu()
    assert got.shape == want.shape == (2, 10, 200, 240, 128)
    assert torch.equal(got != 0, want != 0) or float((got - want).abs().max()) < 1e-4
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-5)
