"""GPU parity of the training-mode forward (batch-statistics BatchNorm, model.train(True) at train.py:148)
against the CPU oracle: outputs and the updated running statistics."""
import numpy as np
import pytest
import torch

from oracle import model_torch as OM
from tests.helpers import load_state, oracle_batch, rel_max_err, synth_frames

pytestmark = pytest.mark.gpu


def test_feature_net_training_forward_matches_oracle(cuda_device):
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(21)
    frames = [f[::3].copy() for f in synth_frames(2, 130)]
    frames[1] = np.concatenate([frames[1], np.tile(np.array([[20.05, 1.05, -1.0, 0.5]], np.float32), (80, 1))])  # > T points
    _, feats, _, coords = oracle_batch(frames, "Car")
    feature = torch.from_numpy(np.concatenate(feats))
    new = {}
    want = OM.svfe_voxelwise(feature, sd, training=True, new_stats=new)
    fnet = vn.FeatureLearningNet("Car")
    load_state(fnet, sd, "")
    fnet = fnet.to(cuda_device).train()
    got = fnet.voxelwise(feature.to(cuda_device))
    assert torch.allclose(got.cpu(), want, rtol=1e-4, atol=1e-5), float((got.cpu() - want).abs().max())
    for name, bn in (("vfe_1.bn.", fnet.vfe_1.bn), ("vfe_2.bn.", fnet.vfe_2.bn)):
        assert torch.allclose(bn.running_mean.cpu(), new[name + "running_mean"], rtol=1e-4, atol=1e-6), name
        assert torch.allclose(bn.running_var.cpu(), new[name + "running_var"], rtol=1e-4, atol=1e-6), name
        assert int(bn.num_batches_tracked) == 1
    # eval afterwards uses the updated running statistics
    sd2 = dict(sd)
    sd2.update(new)
    got_eval = fnet.eval().voxelwise(feature.to(cuda_device))
    assert torch.allclose(got_eval.cpu(), OM.svfe_voxelwise(feature, sd2), rtol=1e-4, atol=1e-5)
    # the drop-in forward (lists in, dense grid out) in training mode
    fnet2 = vn.FeatureLearningNet("Car")
    load_state(fnet2, sd, "")
    fnet2 = fnet2.to(cuda_device).train()
    dense = fnet2([torch.from_numpy(f).to(cuda_device) for f in feats], [torch.from_numpy(c).to(cuda_device) for c in coords])
    want_dense = OM.scatter_dense(want, torch.from_numpy(np.concatenate(coords)), 2, "Car")
    assert torch.allclose(dense.cpu(), want_dense, rtol=1e-4, atol=1e-5)


def test_middle_net_training_forward_matches_oracle(cuda_device):
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(22)
    g = torch.Generator().manual_seed(5)
    B, D, H, W = 2, 10, 40, 32
    dense = torch.zeros((B, D, H, W, 128))
    occ = torch.rand((B, D, H, W), generator=g) < 0.02
    dense[occ] = torch.randn((int(occ.sum()), 128), generator=g)
    new = {}
    want = OM.middle_layers(dense, sd, training=True, new_stats=new)
    mnet = vn.MiddleConvNet()
    load_state(mnet, sd, "")
    mnet = mnet.to(cuda_device).train()
    got = mnet(dense.to(cuda_device))
    assert got.shape == want.shape
    assert rel_max_err(got.cpu(), want) < 1e-2
    for i, layer in enumerate(mnet.middle_layer):
        pre = "middle_layer.%d.batch_norm." % i
        assert torch.allclose(layer.batch_norm.running_mean.cpu(), new[pre + "running_mean"], rtol=2e-2, atol=2e-3), pre
        assert torch.allclose(layer.batch_norm.running_var.cpu(), new[pre + "running_var"], rtol=2e-2, atol=2e-3), pre
    # drop-in ConvMD.forward (NCDHW fp32 in / out) in training mode
    layer = vn.ConvMD(3, 128, 64, 3, (2, 1, 1), (1, 1, 1))
    load_state(layer, {k[len("middle_layer.0."):]: v for k, v in sd.items() if k.startswith("middle_layer.0.")}, "")
    layer = layer.to(cuda_device).train()
    y = layer(dense.permute(0, 4, 1, 2, 3).to(cuda_device))
    ref = OM.middle_layers(dense, sd, upto=1, training=True)
    assert rel_max_err(y.cpu(), ref) < 1e-2
