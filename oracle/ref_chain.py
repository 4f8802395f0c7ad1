"""TEST / BASELINE INFRASTRUCTURE - the UNMODIFIED reference's CPU path for the hot path, timed.

The call sequence of voxelnet/predict.py:52-60 restricted to the path: `pcl_to_voxels(pcl, target, False)`
(voxelnet/utils.py:10-100) -> `prepare_voxel` (voxelnet/dataset.py:101-119) -> `torch.from_numpy` (dataset.py:91-93)
-> `FeatureLearningNet.forward` (voxelnet/model.py:91-108) -> `MiddleConvNet.middle_layer` on the permuted grid +
reshape (model.py:257-262).  The modules are the reference's own classes imported by `oracle/ref_import.py` from
`/root/reference` or from the byte-for-byte copy in `baseline/_ref/` (see `oracle/install_ref.py`); parameters are
loaded by name from an oracle-style state dict.  Same return contract as `oracle.cpu_chain.front_end_cpu`.

Used by `bench.py --impl reference` / `cpu_baseline` (kind "reference") and by tests; never by the product.
"""
from __future__ import annotations

import contextlib
import time
import warnings

import numpy as np
import torch

from . import ref_import as R

_MODULES = {}


def available() -> bool:
    return R.reference_available()


def _modules(sd: dict, target: str):
    """The reference's FeatureLearningNet / MiddleConvNet in eval mode with `sd` loaded (cached per state dict)."""
    key = (id(sd), target)
    if key not in _MODULES:
        ref = R.load_reference()
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            fnet = ref.model.FeatureLearningNet().eval()
            mid = ref.model.MiddleConvNet().eval()
        fnet.load_state_dict({k: v for k, v in sd.items() if k.startswith("vfe_")})
        mid.load_state_dict({k: v for k, v in sd.items() if k.startswith("middle_layer.")}, strict=False)
        _MODULES.clear()
        _MODULES[key] = (fnet, mid)
    return _MODULES[key]


def front_end_reference(frames: list, sd: dict, target: str = "Car", shuffle: bool = False):
    """Returns (output (B,128,H,W) f32, {stage: seconds}).  shuffle=False disables the reference's in-place
    `np.random.shuffle` (utils.py:35) so that the result is comparable point for point; True leaves the reference
    exactly as it runs (the frames are copied first either way, the function mutates its argument)."""
    ref = R.load_reference()
    fnet, mid = _modules(sd, target)
    t = {}
    with warnings.catch_warnings(), (contextlib.nullcontext() if shuffle else R.no_shuffle()), R.ref_model_cfg(target):
        warnings.simplefilter("ignore")
        t0 = time.perf_counter()
        dicts = [ref.utils.pcl_to_voxels(np.array(f, dtype=np.float32, copy=True), target, False) for f in frames]
        feats, _, coords = ref.dataset.prepare_voxel(dicts)
        vf = [torch.from_numpy(f) for f in feats]
        vc = [torch.from_numpy(c) for c in coords]
        t["voxelize"] = time.perf_counter() - t0
        with torch.no_grad():
            t0 = time.perf_counter()
            dense = fnet(vf, vc)
            t["svfe_scatter"] = time.perf_counter() - t0
            t0 = time.perf_counter()
            B, _, H, W, _ = dense.shape
            out = mid.middle_layer(dense.permute(0, 4, 1, 2, 3)).reshape((B, -1, H, W))
            t["conv_middle"] = time.perf_counter() - t0
    t["total"] = sum(t.values())
    return out, t
