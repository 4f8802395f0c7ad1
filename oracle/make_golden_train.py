"""Generate tests/golden/train_grads_ped.npz by running the UNMODIFIED reference in TRAINING mode
(container only; the generating script for the committed fixture).

    python -m oracle.make_golden_train

What runs: the reference's `FeatureLearningNet` (voxelnet/model.py:85-108) and `MiddleConvNet.middle_layer` +
reshape (model.py:206-210, 257-262) under `.train()` - batch-statistics BatchNorm exactly as train.py:148 sets
it - on the Pedestrian grid, fed the reference voxelization of the decimated real KITTI frame stored in
tests/golden/voxelize_kitti0_sub_car.npz, then `.backward()` from a seeded upstream gradient of the
(1,128,200,240) tensor (train.py:149).  Stored: the output's digest-sized summary and the gradient of every
hot-path parameter (VFE and BatchNorm gradients in full, the three conv weight gradients subsampled 1:7 plus
their norms), all float32.
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import model_torch as OM  # noqa: E402
from oracle import ref_import as R  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
STATE_SEED, DOUT_SEED, TARGET, STRIDE = 41, 42, "Pedestrian", 7


def upstream_gradient(shape, seed: int) -> torch.Tensor:
    """Spatially correlated upstream gradient (a loss gradient is not white noise): 9x9 box-filtered randn."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(shape, generator=g)
    return F.avg_pool2d(x, 9, stride=1, padding=4) * 3.0


def main() -> None:
    warnings.filterwarnings("ignore")
    points = np.load(os.path.join(GOLDEN, "voxelize_kitti0_sub_car.npz"))["points"]
    sd = OM.random_state_dict(STATE_SEED)
    with R.ref_model_cfg(TARGET) as ref:
        vox = R.ref_pcl_to_voxels(points, TARGET)
        feat = torch.from_numpy(vox["feature_buffer"])
        coord = torch.from_numpy(np.pad(vox["coordinate_buffer"], ((0, 0), (1, 0)), constant_values=0))
        fnet = ref.model.FeatureLearningNet().train()
        fnet.load_state_dict({k: v for k, v in sd.items() if k.startswith("vfe_")})
        mid = ref.model.MiddleConvNet().train()
        mid.load_state_dict({k: v for k, v in sd.items() if k.startswith("middle_layer.")}, strict=False)
        dense = fnet([feat], [coord])
        B, _, H, W, _ = dense.shape
        y = mid.middle_layer(dense.permute(0, 4, 1, 2, 3)).reshape((B, -1, H, W))
        dout = upstream_gradient(y.shape, DOUT_SEED)
        y.backward(dout)
    out = {
        "state_seed": np.array(STATE_SEED), "dout_seed": np.array(DOUT_SEED), "target": np.array(TARGET),
        "conv_stride": np.array(STRIDE), "n_voxels": np.array(feat.shape[0]),
        "out_abs_max": np.array(float(y.detach().abs().max())), "out_mean": np.array(float(y.detach().mean())),
        "out_sample": y.detach()[0, ::16, ::8, ::8].numpy().copy(),
    }
    named = [("", fnet), ("", mid)]
    for _, mod in named:
        for name, p in mod.named_parameters():
            if not (name.startswith("vfe_") or name.startswith("middle_layer.")) or p.grad is None:
                continue
            g = p.grad.detach().float()
            if name.endswith("conv.weight"):
                out["grad:" + name] = g.reshape(-1)[::STRIDE].numpy().copy()
                out["norm:" + name] = np.array(float(g.norm()))
            else:
                out["grad:" + name] = g.numpy().copy()
    path = os.path.join(GOLDEN, "train_grads_ped.npz")
    np.savez_compressed(path, **out)
    print("%s %.1f KB, K=%d" % (path, os.path.getsize(path) / 1024, feat.shape[0]))
    for k in sorted(out):
        if k.startswith("grad:"):
            print("  %-44s max|g| %.4g" % (k, float(np.abs(out[k]).max())))


if __name__ == "__main__":
    main()
