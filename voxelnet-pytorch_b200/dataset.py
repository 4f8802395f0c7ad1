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


# ---------------------------------------------------------------------------------------------
# Raw-point input pipeline: .bin -> pinned host -> device CSR.
# The reference voxelizes in DataLoader workers and ships 12.85 MB of features per frame to the
# GPU (dataset.py:51-58, model.py:302-303); here only the 1.95 MB of raw points travel.
# ---------------------------------------------------------------------------------------------
def _pinned(t: torch.Tensor) -> torch.Tensor:
    """Page-locked copy when a CUDA device exists (asynchronous H2D needs it); plain memory otherwise, so that
    the packing logic can be exercised on a CPU-only host."""
    return t.pin_memory() if torch.cuda.is_available() else t


def pinned_frame(n_points: int) -> np.ndarray:
    """An (n_points, 4) float32 numpy array in page-locked host memory (plain memory on a host without CUDA): what
    `HostFrameStream.upload_frames` copies to the device without any host-side packing.  The array keeps its
    allocation alive."""
    return _pinned(torch.empty((int(n_points), 4), dtype=torch.float32)).numpy()


def load_velodyne_bin(path: str, pinned: bool = False) -> np.ndarray:
    """dataset.py:51: a KITTI velodyne file -> (N,4) float32 (x, y, z, reflectance).  pinned=True reads the file
    straight into page-locked memory (see pinned_frame)."""
    if not pinned:
        pcl = np.fromfile(path, dtype=np.float32)
        if pcl.size % 4 != 0:
            raise ValueError("%s: %d floats is not a whole number of (x, y, z, r) points" % (path, pcl.size))
        return pcl.reshape(-1, 4)
    import os

    size = os.path.getsize(path)
    if size % 16 != 0:
        raise ValueError("%s: %d bytes is not a whole number of (x, y, z, r) float32 points" % (path, size))
    out = pinned_frame(size // 16)
    with open(path, "rb") as f:
        got = f.readinto(memoryview(out).cast("B"))
    if got != size:
        raise IOError("%s: short read (%d of %d bytes)" % (path, got, size))
    return out


class PinnedFrameBatch:
    """A batch of frames concatenated in pinned host memory with its CSR frame offsets - what
    `HostFrameStream.upload` copies to the device asynchronously.  Buffers are reusable: `fill` packs a
    new list of frames into the same pinned allocation (capacity = `max_points`).

    Packing is done by the library's host thread pool (`vn_host_pack_submit`, include/voxelnet_b200.h): 31 MB per
    Car batch of 16 is more than one thread copies within a 1.5 ms step.  `fill_async` returns at once and the
    copy runs beside the caller; `upload` (or `wait_filled`) waits for it.

    Reuse is safe: `upload` leaves an event on the batch that marks the end of its host->device copy and
    `fill` / `fill_async` wait for that event before they overwrite the pinned points.  To keep the host from
    blocking there, rotate at least (front ends in flight + 2) batches."""

    def __init__(self, n_frames: int, max_points: int):
        self.n_frames = int(n_frames)
        self.points = _pinned(torch.empty((int(max_points), 4), dtype=torch.float32))
        self.offsets = _pinned(torch.zeros(self.n_frames + 1, dtype=torch.int32))
        self.n_points = 0
        self.max_points_per_frame = 0
        self._ticket = 0        # pending vn_host_pack job (0: none)
        self._keep = None       # the frame arrays of the pending job (they must outlive it)
        self._in_flight = None  # torch.cuda.Event: end of the last H2D copy that reads this batch

    def _check(self, frames: list) -> list:
        if len(frames) != self.n_frames:
            raise ValueError("expected %d frames, got %d" % (self.n_frames, len(frames)))
        out = []
        for i, f in enumerate(frames):
            if f.ndim != 2 or f.shape[1] != 4:
                raise ValueError("frame %d has shape %s, expected (N,4)" % (i, f.shape))
            if f.dtype != np.float32 or not f.flags["C_CONTIGUOUS"]:
                f = np.ascontiguousarray(f, dtype=np.float32)  # the reference's loader yields float32 already
            out.append(f)
        total = sum(int(f.shape[0]) for f in out)
        if total > self.points.shape[0]:
            raise ValueError("batch of %d points exceeds the pinned capacity %d" % (total, self.points.shape[0]))
        return out

    def fill_async(self, frames: list) -> "PinnedFrameBatch":
        """Start packing `frames` (list of (N_i,4) float32 arrays); returns immediately."""
        import ctypes as C

        from . import _cabi

        self.wait_filled()
        if self._in_flight is not None:  # the previous upload of this buffer must have left the host memory
            self._in_flight.synchronize()
            self._in_flight = None
        fr = self._check(frames)
        n = self.n_frames
        ptrs = (C.c_void_p * max(n, 1))(*[f.ctypes.data for f in fr])
        sizes = (C.c_int64 * max(n, 1))(*[int(f.shape[0]) for f in fr])
        t = _cabi.lib().vn_host_pack_submit(ptrs, sizes, n, self.points.data_ptr(), self.points.shape[0], self.offsets.data_ptr())
        if t <= 0:
            _cabi.check(int(t), "vn_host_pack_submit")
        self._ticket, self._keep = int(t), (fr, ptrs, sizes)
        self.n_points = sum(int(f.shape[0]) for f in fr)
        self.max_points_per_frame = max([int(f.shape[0]) for f in fr], default=0)
        return self

    def wait_filled(self) -> "PinnedFrameBatch":
        if self._ticket:
            from . import _cabi

            _cabi.check(_cabi.lib().vn_host_pack_wait(self._ticket), "vn_host_pack_wait")
            self._ticket, self._keep = 0, None
        return self

    def fill(self, frames: list) -> "PinnedFrameBatch":
        return self.fill_async(frames).wait_filled()

    @classmethod
    def from_frames(cls, frames: list) -> "PinnedFrameBatch":
        return cls(len(frames), max(sum(int(f.shape[0]) for f in frames), 1)).fill(frames)

    @classmethod
    def from_files(cls, paths: list) -> "PinnedFrameBatch":
        return cls.from_frames([load_velodyne_bin(p) for p in paths])
