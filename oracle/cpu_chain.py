"""TEST INFRASTRUCTURE - the reference's whole CPU path for the hot path, restated end to end.

points -> pcl_to_voxels (voxelnet/utils.py:10-100) -> prepare_voxel (voxelnet/dataset.py:101-119)
-> FeatureLearningNet (voxelnet/model.py:91-108) -> middle_layer + reshape (model.py:257-262),
i.e. the call sequence of voxelnet/predict.py:52-60, with per-stage wall times.

Used by `bench.py`'s `cpu_baseline` leg and `--impl reference` arm (kind "port": the Python reference
cannot travel to the GPU box) and by tests as the end-to-end checker.  The
voxelizer timed here is `pcl_to_voxels_loop`, which keeps the reference's cost structure
(per-point Python loop over a dict); the vectorised `pcl_to_voxels_fast` is only the checker.
"""
from __future__ import annotations

import time

import torch

from . import model_torch as OM
from . import voxelize_np as OV


def front_end_cpu(frames: list, sd: dict, target: str = "Car", faithful_voxelizer: bool = True):
    """Returns (output (B,128,H,W) f32, {stage: seconds})."""
    t = {}
    t0 = time.perf_counter()
    vox = OV.pcl_to_voxels_loop if faithful_voxelizer else OV.pcl_to_voxels_fast
    dicts = [vox(f, target) for f in frames]
    feats, _, coords = OV.prepare_voxel(dicts)
    t["voxelize"] = time.perf_counter() - t0
    with torch.no_grad():
        t0 = time.perf_counter()
        dense = OM.feature_learning_net([torch.from_numpy(f) for f in feats], [torch.from_numpy(c) for c in coords], sd, target)
        t["svfe_scatter"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        out = OM.middle_layers(dense, sd)
        t["conv_middle"] = time.perf_counter() - t0
    t["total"] = sum(t.values())
    return out, t
