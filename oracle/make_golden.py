"""Generate tests/golden/*.npz by running the UNMODIFIED reference (container only).

    python -m oracle.make_golden

The reference ships no golden vectors, so the pin is the reference's own code
executed here: `pcl_to_voxels` (voxelnet/utils.py:10-100, shuffle disabled), `FeatureLearningNet`,
`VFELayer` (voxelnet/model.py:60-108) and `MiddleConvNet.middle_layer` + reshape
(model.py:206-210,257-262), imported from /root/reference through oracle/ref_import.py.
Inputs are either a decimated real KITTI frame from the reference's bundled drive (stored in
the fixture) or regenerated from seeds at test time; parameters come from
`oracle.model_torch.random_state_dict(seed)` loaded into the reference modules by name.
"""
from __future__ import annotations

import hashlib
import os
import sys
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import model_torch as OM  # noqa: E402
from oracle import ref_import as R  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def voxel_fixture(pcl: np.ndarray, target: str, keep_points: bool) -> dict:
    v = R.ref_pcl_to_voxels(pcl, target)
    out = {
        "target": np.array(target),
        "coordinate_buffer": v["coordinate_buffer"].astype(np.int16),
        "number_buffer": v["number_buffer"].astype(np.uint8),
        "feature_sha256": np.array(sha(v["feature_buffer"])),
        "coordinate_sha256": np.array(sha(v["coordinate_buffer"])),
        "number_sha256": np.array(sha(v["number_buffer"])),
        "feature_head": v["feature_buffer"][:48].copy(),
    }
    if keep_points:
        out["points"] = pcl
    return out


def synth_frame(seed: int, **kw) -> np.ndarray:
    import voxelnet_pytorch_b200.synth as synth

    return synth.lidar_frame(seed, **kw)


def main() -> None:
    warnings.filterwarnings("ignore")
    os.makedirs(GOLDEN, exist_ok=True)
    ref = R.load_reference()

    # ---- voxelizer ---------------------------------------------------------------------------
    kitti = np.fromfile(R.kitti_frame_path(0), dtype=np.float32).reshape(-1, 4)
    sub = np.ascontiguousarray(kitti[::6])  # ~20k real points keep the fixture small
    np.savez_compressed(os.path.join(GOLDEN, "voxelize_kitti0_sub_car.npz"), **voxel_fixture(sub, "Car", True))
    np.savez_compressed(os.path.join(GOLDEN, "voxelize_kitti0_sub_ped.npz"), **voxel_fixture(sub, "Pedestrian", False))
    # full-size frames: digests only (inputs regenerated from the seed / re-read from the fixture)
    for seed, target in ((11, "Car"), (12, "Pedestrian")):
        f = voxel_fixture(synth_frame(seed), target, False)
        f["seed"] = np.array(seed)
        np.savez_compressed(os.path.join(GOLDEN, "voxelize_synth%d_%s.npz" % (seed, target.lower()[:3])), **f)
    # a permuted frame: selection depends on array order
    perm = np.random.default_rng(5).permutation(len(sub))
    np.savez_compressed(os.path.join(GOLDEN, "voxelize_kitti0_sub_perm_car.npz"),
                        **voxel_fixture(np.ascontiguousarray(sub[perm]), "Car", False), perm_seed=np.array(5))

    # ---- model: reference modules with oracle-generated parameters ----------------------------------
    sd = OM.random_state_dict(21)
    vox = R.ref_pcl_to_voxels(sub, "Car")
    K = 320
    feat = torch.from_numpy(vox["feature_buffer"][:K].copy())
    coord = torch.from_numpy(np.pad(vox["coordinate_buffer"][:K], ((0, 0), (1, 0)), constant_values=0))
    with torch.no_grad():
        fnet = ref.model.FeatureLearningNet().eval()
        fnet.load_state_dict({k: v for k, v in sd.items() if k.startswith("vfe_")})
        dense = fnet([feat], [coord])
        voxelwise = dense[coord[:, 0], coord[:, 1], coord[:, 2], coord[:, 3]].numpy()
        nnz_rows = int((dense.abs().sum(-1) != 0).sum())
        mask = feat.max(dim=2, keepdim=True)[0] != 0
        vfe1 = fnet.vfe_1(feat, mask).numpy()
        g = torch.Generator().manual_seed(22)
        x2 = torch.randn((64, 35, 32), generator=g)
        m2 = torch.rand((64, 35, 1), generator=g) > 0.3
        vfe2 = fnet.vfe_2(x2, m2).numpy()

        mid = ref.model.MiddleConvNet().eval()
        mid.load_state_dict({k: v for k, v in sd.items() if k.startswith("middle_layer.")}, strict=False)
        g = torch.Generator().manual_seed(23)
        B, D, H, W = 1, 10, 24, 16
        xd = torch.randn((B, D, H, W, 128), generator=g) * (torch.rand((B, D, H, W, 1), generator=g) > 0.8)
        y = mid.middle_layer(xd.permute(0, 4, 1, 2, 3))
        y1 = mid.middle_layer[0](xd.permute(0, 4, 1, 2, 3))
        mid_out = y.reshape((B, -1, H, W)).numpy()
    np.savez_compressed(
        os.path.join(GOLDEN, "model_ref_outputs.npz"),
        state_seed=np.array(21), K=np.array(K), voxelwise=voxelwise, dense_nonzero_rows=np.array(nnz_rows),
        vfe1_out=vfe1, vfe2_seed=np.array(22), vfe2_out=vfe2, middle_seed=np.array(23),
        middle_shape=np.array([B, D, H, W]), middle_out=mid_out, middle_layer0_out=y1.numpy(),
    )
    for f in sorted(os.listdir(GOLDEN)):
        print("%-44s %8.1f KB" % (f, os.path.getsize(os.path.join(GOLDEN, f)) / 1024))


if __name__ == "__main__":
    main()
