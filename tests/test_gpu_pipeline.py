"""GPU parity: the fused front end (points -> (B,128,H,W)) vs the oracle chain, and state reuse."""
import numpy as np
import pytest
import torch

from oracle import model_torch as OM
from tests.helpers import load_state, oracle_batch, rel_max_err, synth_frames

pytestmark = pytest.mark.gpu


def _build(cuda_device, target, sd):
    import voxelnet_pytorch_b200 as vn

    fnet = vn.FeatureLearningNet(target).eval()
    mnet = vn.MiddleConvNet().eval()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    return vn.VoxelNetFrontEnd(fnet.to(cuda_device), mnet.to(cuda_device), target)


def test_front_end_matches_oracle_ped(cuda_device):
    """Full chain on the Pedestrian grid (smaller: the CPU oracle conv finishes in seconds)."""
    sd = OM.random_state_dict(11)
    frames = synth_frames(2, 30)
    fe = _build(cuda_device, "Pedestrian", sd)
    got = fe.forward_frames(frames).cpu()
    offs = fe.voxel_counts()
    _, feats, _, coords = oracle_batch(frames, "Pedestrian")
    assert list(np.diff(offs)) == [len(f) for f in feats]
    dense = OM.feature_learning_net([torch.from_numpy(f) for f in feats], [torch.from_numpy(c) for c in coords], sd, "Pedestrian")
    want = OM.middle_layers(dense, sd)
    assert got.shape == want.shape == (2, 128, 200, 240)
    assert rel_max_err(got, want) < 1e-2
    # second call with different frames reuses the resident grid: it must have been re-zeroed
    frames2 = synth_frames(2, 40)
    got2 = fe.forward_frames(frames2).cpu()
    _, feats2, _, coords2 = oracle_batch(frames2, "Pedestrian")
    dense2 = OM.feature_learning_net([torch.from_numpy(f) for f in feats2], [torch.from_numpy(c) for c in coords2], sd, "Pedestrian")
    want2 = OM.middle_layers(dense2, sd)
    assert rel_max_err(got2, want2) < 1e-2
    assert int(fe.dense.view(torch.int16).ne(0).sum()) == 0


def test_front_end_car_shapes_and_determinism(cuda_device):
    sd = OM.random_state_dict(12)
    frames = synth_frames(2, 50)
    fe = _build(cuda_device, "Car", sd)
    a = fe.forward_frames(frames).clone()
    b = fe.forward_frames(frames)
    assert a.shape == (2, 128, 400, 352)
    assert torch.equal(a, b)
    assert bool(torch.isfinite(a).all())
