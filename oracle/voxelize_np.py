"""TEST INFRASTRUCTURE - CPU oracle for the voxelizer and the batch packer.

Numpy restatement of the reference's `pcl_to_voxels` (voxelnet/utils.py:10-100) and
`prepare_voxel` (voxelnet/dataset.py:101-119).  Only `tests/`, `__graft_entry__.smoke()` and
`bench.py`'s cpu-baseline / `--impl reference` legs may import this module; the product package
never does.

Parity pinning: the reference ships no golden vectors or tests for this path, so
this oracle is pinned by running the reference itself here: `tests/test_oracle_vs_reference.py`
compares both functions below with the imported, unmodified reference on KITTI and synthetic
frames (container only), and `oracle/make_golden.py` freezes reference outputs into
`tests/golden/` so the pin travels to machines that have no `/root/reference`.

Two restatements are kept:

* `pcl_to_voxels_loop`  - statement-for-statement the reference algorithm, including its
  per-point Python loop over a dict (utils.py:74-84).  This is what `bench.py --impl reference`
  times: it has the reference's cost structure.
* `pcl_to_voxels_fast`  - the same results through a stable sort; used as the checker at full
  sizes (seconds -> milliseconds).  Tested equal to the loop version bit for bit.

The reference's in-place `np.random.shuffle(pcl)` (utils.py:35) is NOT performed: point order is
the caller's.  Callers that want the reference's randomisation permute the input
with their own seeded permutation and hand the same array to both sides.
"""
from __future__ import annotations

import numpy as np

# voxelnet/utils.py:24-33 - all float32, axis order (z, y, x) for voxel_size / grid_size and
# (x, y, z) for lidar_coord.
_CLASS_PARAMS = {
    "Car": dict(voxel_size=(0.4, 0.2, 0.2), grid_size=(10, 400, 352), lidar_coord=(0, 40, 3), T=35),
    "Other": dict(voxel_size=(0.4, 0.2, 0.2), grid_size=(10, 200, 240), lidar_coord=(0, 20, 3), T=45),
}


def class_params(target: str) -> dict:
    """utils.py:24-33: 'Car' selects the Car grid, anything else the Pedestrian/Cyclist grid."""
    p = _CLASS_PARAMS["Car" if target == "Car" else "Other"]
    return dict(
        voxel_size=np.array(p["voxel_size"], dtype=np.float32),
        grid_size=np.array(p["grid_size"], dtype=np.float32),
        lidar_coord=np.array(p["lidar_coord"], dtype=np.float32),
        T=p["T"],
    )


def _index_and_crop(pcl: np.ndarray, target: str):
    """utils.py:37-61: fp32 shift, fp32 IEEE divide, floor, int64 cast, range crop."""
    p = class_params(target)
    shifted_coord = pcl[:, :3] + p["lidar_coord"]
    voxel_index = np.floor(shifted_coord[:, ::-1] / p["voxel_size"]).astype(int)
    grid = p["grid_size"]
    keep = np.ones(len(pcl), dtype=bool)
    for axis in range(3):
        keep &= (voxel_index[:, axis] >= 0) & (voxel_index[:, axis] < grid[axis])
    return pcl[keep], voxel_index[keep], p


def _centroid_offsets(feature_buffer: np.ndarray, number_buffer: np.ndarray) -> None:
    """utils.py:87-88: offsets for ALL T slots; fp32 sum over t, / int64 -> float64, store fp32."""
    K = len(number_buffer)
    feature_buffer[:, :, -3:] = feature_buffer[:, :, :3] - feature_buffer[:, :, :3].sum(
        axis=1, keepdims=True
    ) / number_buffer.reshape(K, 1, 1)


def pcl_to_voxels_loop(pcl: np.ndarray, target: str) -> dict:
    """Statement-level restatement of utils.py:37-94 (shuffle at :35 omitted)."""
    pcl, voxel_index, p = _index_and_crop(pcl, target)
    T = p["T"]
    coordinate_buffer = np.unique(voxel_index, axis=0)  # utils.py:63
    K = len(coordinate_buffer)
    number_buffer = np.zeros(shape=(K), dtype=np.int64)
    feature_buffer = np.zeros(shape=(K, T, 7), dtype=np.float32)
    slot_of = {}
    for row in range(K):
        slot_of[tuple(coordinate_buffer[row])] = row
    for cell, point in zip(voxel_index, pcl):  # utils.py:78-84
        row = slot_of[tuple(cell)]
        n = number_buffer[row]
        if n < T:
            feature_buffer[row, n, :4] = point
            number_buffer[row] += 1
    _centroid_offsets(feature_buffer, number_buffer)
    return {
        "feature_buffer": feature_buffer,
        "coordinate_buffer": coordinate_buffer,
        "number_buffer": number_buffer,
    }


def pcl_to_voxels_fast(pcl: np.ndarray, target: str) -> dict:
    """Same outputs as `pcl_to_voxels_loop`, via a stable sort on the linear cell key.

    np.unique(axis=0) orders rows lexicographically by (z, y, x) (utils.py:63), i.e. ascending
    key = (z*H + y)*W + x; "first T points in array order" (utils.py:78-84) is
    the first T entries of each key's run after a STABLE sort.
    """
    pcl, voxel_index, p = _index_and_crop(pcl, target)
    T = p["T"]
    _, H, W = (int(g) for g in p["grid_size"])
    key = (voxel_index[:, 0] * H + voxel_index[:, 1]) * W + voxel_index[:, 2]
    order = np.argsort(key, kind="stable")
    skey = key[order]
    if len(skey):
        first = np.flatnonzero(np.concatenate(([True], skey[1:] != skey[:-1])))
    else:
        first = np.zeros((0,), dtype=np.int64)
    K = len(first)
    ukey = skey[first]
    coordinate_buffer = np.stack([ukey // (H * W), (ukey // W) % H, ukey % W], axis=1).astype(np.int64)
    coordinate_buffer = coordinate_buffer.reshape(K, 3)
    run_len = np.diff(np.concatenate((first, [len(skey)])))
    row_of_sorted = np.repeat(np.arange(K), run_len)
    rank = np.arange(len(skey)) - first[row_of_sorted]
    keep = rank < T
    number_buffer = np.minimum(run_len, T).astype(np.int64)
    feature_buffer = np.zeros(shape=(K, T, 7), dtype=np.float32)
    feature_buffer[row_of_sorted[keep], rank[keep], :4] = pcl[order[keep]]
    _centroid_offsets(feature_buffer, number_buffer)
    return {
        "feature_buffer": feature_buffer,
        "coordinate_buffer": coordinate_buffer,
        "number_buffer": number_buffer,
    }


def prepare_voxel(voxels: list) -> tuple:
    """dataset.py:101-119: left-pad each frame's (K,3) coords with its batch index -> (K,4)."""
    features, numbers, coordinates = [], [], []
    for i, v in enumerate(voxels):
        features.append(v["feature_buffer"])
        numbers.append(v["number_buffer"])
        coordinates.append(np.pad(v["coordinate_buffer"], ((0, 0), (1, 0)), mode="constant", constant_values=i))
    return features, numbers, coordinates


def voxel_dict_digest(v: dict) -> dict:
    """Small, order-sensitive fingerprint of a voxel dict (for golden files)."""
    import hashlib

    out = {}
    for k in ("feature_buffer", "coordinate_buffer", "number_buffer"):
        a = np.ascontiguousarray(v[k])
        out[k] = dict(shape=list(a.shape), dtype=str(a.dtype), sha256=hashlib.sha256(a.tobytes()).hexdigest())
    return out
