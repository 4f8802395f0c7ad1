"""Voxelization on the GPU behind the reference's `pcl_to_voxels` interface.

Mirror of voxelnet/utils.py:10-100 (same name, arguments, returned dict and side effect on
`pcl`), plus the batched device-resident form the fast pipeline uses.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _cabi, ops
from .config import grid_config


def _device(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError(
            "voxelnet_pytorch_b200 needs a CUDA device: voxelization runs as sm_100a kernels and there is no CPU fallback"
        )
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


class VoxelBatch:
    """Device-resident result of voxelizing a batch of frames.

    feature (cap,T,7) f32 | coordinate (cap,4) i64 [b,z,y,x] | number (cap,) i64 |
    voxel_offsets (B+1,) i32 exclusive prefix of voxels per frame (device); rows past
    voxel_offsets[B] are unspecified.
    """

    __slots__ = ("feature", "coordinate", "number", "voxel_offsets", "n_frames")

    def __init__(self, feature, coordinate, number, voxel_offsets, n_frames):
        self.feature, self.coordinate, self.number = feature, coordinate, number
        self.voxel_offsets, self.n_frames = voxel_offsets, n_frames

    @property
    def n_voxels_dev(self) -> torch.Tensor:
        """1-element int32 view holding the live total K (stays on the device)."""
        return self.voxel_offsets[self.n_frames :]


def voxelize_batch(points, frame_offsets, max_points_per_frame: int, target: str = "Car", capacity=None,
                   out: VoxelBatch = None, workspace=None) -> VoxelBatch:
    """Voxelize concatenated frames that already live on the device.

    points (N,4) f32 cuda, frame_offsets (B+1,) i32 cuda.  With `capacity=None` the exact number
    of voxels is read back (one host sync) and outputs are sized exactly, as the reference's
    arrays are; with an integer capacity nothing synchronises and outputs have `capacity` rows.
    """
    grid = _cabi.grid_for_target(target)
    dev = points.device
    n_frames = frame_offsets.numel() - 1
    if workspace is None:
        nbytes = ops.voxelize_workspace_bytes(grid, n_frames, points.shape[0])
        workspace = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    voxel_offsets = out.voxel_offsets if out is not None else torch.empty(n_frames + 1, dtype=torch.int32, device=dev)
    ops.voxelize_index(grid, points, frame_offsets, max_points_per_frame, voxel_offsets, workspace)
    if out is None:
        if capacity is None:
            capacity = ops.voxel_offsets_to_host(voxel_offsets)[-1]
        T = grid.T
        out = VoxelBatch(
            torch.empty((capacity, T, 7), dtype=torch.float32, device=dev),
            torch.empty((capacity, 4), dtype=torch.int64, device=dev),
            torch.empty((capacity,), dtype=torch.int64, device=dev),
            voxel_offsets,
            n_frames,
        )
    ops.voxelize_fill(grid, points, frame_offsets, max_points_per_frame, voxel_offsets, out.feature, out.coordinate,
                      out.number, workspace)
    return out


def pcl_to_voxels(pcl: np.ndarray, target: str, verbose: bool = False, shuffle: bool = True, device=None) -> dict:
    """Drop-in for the reference's `pcl_to_voxels` (voxelnet/utils.py:10-100).

    pcl (N,4) float32 [x,y,z,r]; target 'Car' or anything else (Pedestrian/Cyclist grid).
    Returns {'feature_buffer' (K,T,7) f32, 'coordinate_buffer' (K,3) int64 (z,y,x) in
    lexicographic order, 'number_buffer' (K,) int64}.

    Like the reference (utils.py:35) the caller's array is shuffled IN PLACE with
    `np.random.shuffle` first - same global RNG, same call, so a seeded run selects the same
    points as the reference does; `shuffle=False` keeps the caller's order.
    """
    dev = _device(device)
    if shuffle:
        np.random.shuffle(pcl)
    pts = np.ascontiguousarray(pcl, dtype=np.float32)
    if pts.ndim != 2 or pts.shape[1] != 4:
        raise ValueError("pcl must have shape (N, 4), got %s" % (pts.shape,))
    grid = grid_config(target)
    n = pts.shape[0]
    if n == 0:
        vb_feature = np.zeros((0, grid.points_per_voxel, 7), np.float32)
        vb_coord, vb_num = np.zeros((0, 3), np.int64), np.zeros((0,), np.int64)
    else:
        with torch.cuda.device(dev):
            points = torch.from_numpy(pts).to(dev)
            offsets = torch.tensor([0, n], dtype=torch.int32, device=dev)
            vb = voxelize_batch(points, offsets, n, target)
            vb_feature = vb.feature.cpu().numpy()
            vb_coord = vb.coordinate[:, 1:].contiguous().cpu().numpy()
            vb_num = vb.number.cpu().numpy()
    voxel_dict = {
        "feature_buffer": vb_feature,
        "coordinate_buffer": vb_coord,
        "number_buffer": vb_num,
    }
    if verbose:  # utils.py:95-98
        print(f"Coordinate buffer shape: {voxel_dict['coordinate_buffer'].shape}")
        print(f"Feature buffer shape: {voxel_dict['feature_buffer'].shape}")
        print(f"Number buffer shape: {voxel_dict['number_buffer'].shape}")
    return voxel_dict


def pack_frames(frames: list, device=None, pin: bool = False):
    """Concatenate host frames [(N_i,4) f32] -> (points cuda (N,4), frame_offsets cuda (B+1,) i32,
    max points per frame).  `pin=True` stages through pinned memory for an async copy."""
    dev = _device(device)
    sizes = [int(f.shape[0]) for f in frames]
    offs = np.zeros(len(frames) + 1, dtype=np.int32)
    np.cumsum(sizes, out=offs[1:])
    host = torch.from_numpy(np.ascontiguousarray(np.concatenate(frames, axis=0), dtype=np.float32))
    if pin:
        host = host.pin_memory()
    points = host.to(dev, non_blocking=pin)
    frame_offsets = torch.from_numpy(offs).to(dev)
    return points, frame_offsets, max(sizes) if sizes else 0
