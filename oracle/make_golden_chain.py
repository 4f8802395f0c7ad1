"""Generate tests/golden/chain_kitti0_car.npz + tests/golden/kitti_0000000000.bin (container only).

    python -m oracle.make_golden_chain

The WHOLE hot path of the UNMODIFIED reference on one full-size real frame at the Car grid - the call sequence of
voxelnet/predict.py:52-60: pcl_to_voxels (utils.py:10-100, shuffle disabled) -> prepare_voxel (dataset.py:101-119) ->
FeatureLearningNet (model.py:91-108) -> middle_layer + reshape (model.py:257-262) - frozen so that the GPU box, where the
reference may be absent, can hold the CUDA path (and the oracle) to a reference-generated result at the benchmarked size.

The input is frame 0 of the KITTI drive bundled with the reference (121,015 points); it is committed next to the golden
(1.9 MB) so that a real frame reaches the GPU tests.  Stored: digests of the voxel dict, the (1,128,400,352) output
sub-sampled 1:4 in H and W (all channels), its max-abs and per-channel sums over the full tensor.
Parameters: `oracle.model_torch.random_state_dict(41)` loaded into the reference modules by name.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import model_torch as OM  # noqa: E402
from oracle import ref_chain  # noqa: E402
from oracle import ref_import as R  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
STATE_SEED = 41


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main() -> None:
    src = R.kitti_frame_path(0)
    dst = os.path.join(GOLDEN, "kitti_0000000000.bin")
    shutil.copyfile(src, dst)
    pcl = np.fromfile(dst, dtype=np.float32).reshape(-1, 4)
    v = R.ref_pcl_to_voxels(pcl, "Car")
    sd = OM.random_state_dict(STATE_SEED)
    out, _ = ref_chain.front_end_reference([pcl], sd, "Car", shuffle=False)
    out = out.numpy()
    np.savez_compressed(
        os.path.join(GOLDEN, "chain_kitti0_car.npz"),
        state_seed=np.array(STATE_SEED), n_points=np.array(len(pcl)), K=np.array(len(v["number_buffer"])),
        coordinate_sha256=np.array(sha(v["coordinate_buffer"])), number_sha256=np.array(sha(v["number_buffer"])),
        feature_sha256=np.array(sha(v["feature_buffer"])),
        out_shape=np.array(out.shape), out_sample=out[:, :, 1::4, 2::4].copy(), out_max_abs=np.array(np.abs(out).max()),
        out_channel_sums=out.astype(np.float64).sum(axis=(0, 2, 3)),
    )
    for f in ("kitti_0000000000.bin", "chain_kitti0_car.npz"):
        print("%-28s %8.1f KB" % (f, os.path.getsize(os.path.join(GOLDEN, f)) / 1024))
    print("K =", len(v["number_buffer"]), "max|out| =", float(np.abs(out).max()))


if __name__ == "__main__":
    main()
