"""Shared helpers for the parity tests (the oracle is the checker, never the product)."""
import numpy as np
import torch

from oracle import model_torch as OM
from oracle import voxelize_np as OV


def synth_frames(n, seed0=0, **kw):
    from voxelnet_pytorch_b200 import synth

    return synth.lidar_batch(n, seed0, **kw)


def oracle_batch(frames, target):
    """Oracle voxelization of a list of frames -> (dicts, features, numbers, coordinates[b,z,y,x])."""
    dicts = [OV.pcl_to_voxels_fast(f, target) for f in frames]
    feats, nums, coords = OV.prepare_voxel(dicts)
    return dicts, feats, nums, coords


def assert_voxels_equal(got: dict, want: dict):
    """Bit-exact comparison of two voxel dicts (coords, counts, and every feature bit)."""
    for k in ("coordinate_buffer", "number_buffer", "feature_buffer"):
        g, w = np.asarray(got[k]), np.asarray(want[k])
        assert g.shape == w.shape, (k, g.shape, w.shape)
        assert g.dtype == w.dtype, (k, g.dtype, w.dtype)
    assert np.array_equal(got["coordinate_buffer"], want["coordinate_buffer"])
    assert np.array_equal(got["number_buffer"], want["number_buffer"])
    g = np.ascontiguousarray(got["feature_buffer"]).view(np.uint32)
    w = np.ascontiguousarray(want["feature_buffer"]).view(np.uint32)
    bad = g != w
    assert not bad.any(), "feature_buffer differs in %d of %d words" % (bad.sum(), bad.size)


def load_state(module, sd, prefix):
    """Load the oracle-style flat state dict entries under `prefix` into a module."""
    own = set(module.state_dict().keys())
    sub = {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix) and k[len(prefix):] in own}
    missing, unexpected = module.load_state_dict(sub, strict=False)
    assert not unexpected, unexpected
    assert not missing, missing


def rel_max_err(got: torch.Tensor, want: torch.Tensor) -> float:
    """max |got - want| / max |want|  (the conv tolerance metric, SURVEY hard part 5)."""
    return float((got.double() - want.double()).abs().max() / want.double().abs().max().clamp_min(1e-30))
