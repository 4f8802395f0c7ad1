"""Loading the golden fixtures written by oracle/make_golden.py (reference outputs)."""
import hashlib
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def kitti_sub_points():
    return load("voxelize_kitti0_sub_car.npz")["points"]


def voxel_cases():
    """(fixture name, target, points) for every voxelizer golden."""
    from voxelnet_pytorch_b200 import synth

    sub = kitti_sub_points()
    perm = np.random.default_rng(5).permutation(len(sub))
    return [
        ("voxelize_kitti0_sub_car.npz", "Car", sub),
        ("voxelize_kitti0_sub_ped.npz", "Pedestrian", sub),
        ("voxelize_kitti0_sub_perm_car.npz", "Car", np.ascontiguousarray(sub[perm])),
        ("voxelize_synth11_car.npz", "Car", synth.lidar_frame(11)),
        ("voxelize_synth12_ped.npz", "Pedestrian", synth.lidar_frame(12)),
    ]


def check_voxel_dict_against_golden(v, name):
    g = load(name)
    assert np.array_equal(v["coordinate_buffer"], g["coordinate_buffer"].astype(np.int64)), name
    assert np.array_equal(v["number_buffer"], g["number_buffer"].astype(np.int64)), name
    assert v["coordinate_buffer"].dtype == np.int64 and v["number_buffer"].dtype == np.int64
    assert v["feature_buffer"].dtype == np.float32
    assert sha(v["coordinate_buffer"]) == str(g["coordinate_sha256"]), name
    assert sha(v["number_buffer"]) == str(g["number_sha256"]), name
    head = g["feature_head"]
    assert np.array_equal(v["feature_buffer"][: len(head)].view(np.uint32), head.view(np.uint32)), name
    assert sha(v["feature_buffer"]) == str(g["feature_sha256"]), name
