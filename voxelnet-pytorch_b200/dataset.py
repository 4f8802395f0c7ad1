"""Batch packing behind the reference's `prepare_voxel` / `collate_fn` interface.

Mirror of voxelnet/dataset.py:101-119 and :70-97.  These are host-side list/array shuffles
(no arithmetic); the device-side equivalent - the batch index as column 0 of the int64
coordinates - is produced directly by the voxelizer kernel (`utils.voxelize_batch`).
"""
from __future__ import annotations

import numpy as np
import torch


def prepare_voxel(voxels: list) -> tuple:
    """dataset.py:101-119: list of voxel dicts -> (features, numbers, coordinates) lists, where
    coordinates[i] is (K_i, 4) int64 [i, z, y, x]."""
    features, numbers, coordinates = [], [], []
    for i, voxel in enumerate(voxels):
        features.append(voxel["feature_buffer"])
        numbers.append(voxel["number_buffer"])
        zyx = voxel["coordinate_buffer"]
        bzyx = np.empty((zyx.shape[0], 4), dtype=zyx.dtype)
        bzyx[:, 0] = i
        bzyx[:, 1:] = zyx
        coordinates.append(bzyx)
    return features, numbers, coordinates


def collate_fn(parts: tuple) -> tuple:
    """dataset.py:70-97: samples (tag, rgb, raw_lidar, label, voxel_dict) -> the 7-tuple the
    model reads (positions 1, 2, 4: labels, feature tensors, coordinate tensors)."""
    tag = [p[0] for p in parts]
    rgb = [p[1] for p in parts]
    raw_lidar = [p[2] for p in parts]
    label = [p[3] for p in parts]
    voxel_features, voxel_numbers, voxel_coordinates = prepare_voxel([p[4] for p in parts])
    return (
        tag,
        np.array(label, dtype=object),
        [torch.from_numpy(f) for f in voxel_features],
        np.array(voxel_numbers, dtype=object),
        [torch.from_numpy(c) for c in voxel_coordinates],
        np.array(rgb, dtype=object),
        np.array(raw_lidar, dtype=object),
    )
