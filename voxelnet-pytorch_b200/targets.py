"""RPN target assignment - host mirror of the reference's anchor / target functions over the device kernels.

Drop-in names and argument meaning (reference voxelnet/utils.py): `generate_anchors` (:104-130), `label_to_gt_box_3d`
(:176-209), `bbox_iou` (:344-373), `generate_targets` (:376-462; call site voxelnet/model.py:309).  The anchor x
ground-truth part - 70,400 anchors x G boxes per Car frame, a pure-Python double loop in the reference - runs in
`vn_bbox_iou` / `vn_rpn_targets` (csrc/vn_targets.cu); the per-box preparation (label text -> lidar box -> float32
stand-up box, a few dozen boxes per frame) stays on the host in numpy, written so that every value handed to the device
has the bits the reference hands to its own loop.

The reference reads its thresholds from a module-level config fixed to one class at import time (config.py:36); here
the class is the explicit `cls_name` argument, as everywhere in this package.

There is no CPU path: the functions raise without a CUDA device.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import ops


@dataclass(frozen=True)
class AnchorConfig:
    """cfg.OBJECT.* of config.py:38-92 for one class."""
    x_range: tuple
    y_range: tuple
    feature_width: int
    feature_height: int
    anchor_l: float
    anchor_w: float
    anchor_h: float
    anchor_z: float
    rpn_pos_iou: float
    rpn_neg_iou: float


ANCHORS = {
    "Car": AnchorConfig((0, 70.4), (-40, 40), 176, 200, 3.9, 1.6, 1.56, -1.0 - 1.56 / 2, 0.6, 0.45),
    "Pedestrian": AnchorConfig((0, 48), (-20, 20), 120, 100, 0.8, 0.6, 1.73, -0.6 - 1.73 / 2, 0.5, 0.35),
    "Cyclist": AnchorConfig((0, 48), (-20, 20), 120, 100, 1.76, 0.6, 1.73, -0.6 - 1.73 / 2, 0.5, 0.35),
}

# cfg.CALIB.T_VELO_2_CAM / R_RECT_0 (config.py:103-114): the fixed calibration the reference falls back to.
T_VELO_2_CAM = np.array([
    [7.49916597e-03, -9.99971248e-01, -8.65110297e-04, -6.71807577e-03],
    [1.18652889e-02, 9.54520517e-04, -9.99910318e-01, -7.33152811e-02],
    [9.99882833e-01, 7.49141178e-03, 1.18719929e-02, -2.78557062e-01],
    [0, 0, 0, 1],
])
R_RECT_0 = np.array([
    [0.99992475, 0.00975976, -0.00734152, 0],
    [-0.0097913, 0.99994262, -0.00430371, 0],
    [0.00729911, 0.0043753, 0.99996319, 0],
    [0, 0, 0, 1],
])


def anchor_config(cls_name: str) -> AnchorConfig:
    if cls_name not in ANCHORS:
        raise ValueError("unknown class %r (Car, Pedestrian, Cyclist)" % (cls_name,))
    return ANCHORS[cls_name]


def generate_anchors(cls_name: str = "Car") -> np.ndarray:
    """(FEATURE_HEIGHT, FEATURE_WIDTH, 2, 7) float64 anchors (x, y, z, h, w, l, r), r in {0, pi/2} - utils.py:104-130."""
    c = anchor_config(cls_name)
    xs = np.linspace(c.x_range[0], c.x_range[1], c.feature_width)
    ys = np.linspace(c.y_range[0], c.y_range[1], c.feature_height)
    out = np.empty((c.feature_height, c.feature_width, 2, 7), dtype=np.float64)
    out[..., 0] = xs[None, :, None]
    out[..., 1] = ys[:, None, None]
    out[..., 2] = c.anchor_z
    out[..., 3] = c.anchor_h
    out[..., 4] = c.anchor_w
    out[..., 5] = c.anchor_l
    out[..., 0, 6] = 0
    out[..., 1, 6] = 90 / 180 * np.pi
    return out


def _camera_box_to_lidar(box, t_velo_2_cam, r_rect_0) -> list:
    """One (x, y, z, h, w, l, ry) camera box -> lidar box, utils.py:146-173 with the yaw wrap of :133-143."""
    x, y, z, h, w, l, ry = box
    p = np.array([x, y, z, 1])
    p = np.matmul(np.linalg.inv(r_rect_0), p)
    p = np.matmul(np.linalg.inv(t_velo_2_cam), p)
    rz = -ry - np.pi / 2
    while rz >= np.pi / 2:
        rz -= np.pi
    while rz < -np.pi / 2:
        rz += np.pi
    if abs(rz + np.pi / 2) < 5 / 180 * np.pi:
        rz = np.pi / 2
    return [p[0], p[1], p[2], h, w, l, rz]


def label_to_gt_box_3d(labels, cls_name: str, coordinate: str, T_VELO_2_CAM=None, R_RECT_0=None) -> list:
    """KITTI label lines per frame -> list of (G_f, 7) float64 boxes (x, y, z, h, w, l, r) - utils.py:176-209.

    'Car' accepts Car and Van; an unknown class name accepts every line, as in the reference."""
    accepted = {"Car": ("Car", "Van"), "Pedestrian": ("Pedestrian",), "Cyclist": ("Cyclist",)}.get(cls_name, ())
    tv = np.array(globals()["T_VELO_2_CAM"] if T_VELO_2_CAM is None else T_VELO_2_CAM)
    rr = np.array(globals()["R_RECT_0"] if R_RECT_0 is None else R_RECT_0)
    out = []
    for frame_lines in labels:
        boxes = []
        for line in frame_lines:
            fields = line.split()
            if not accepted or fields[0] in accepted:
                h, w, l, x, y, z, r = (float(v) for v in fields[-7:])
                boxes.append(np.array([x, y, z, h, w, l, r]))
        if coordinate == "lidar":
            boxes = [_camera_box_to_lidar(b, tv, rr) for b in boxes]
        out.append(np.array(boxes, dtype=np.float64).reshape(-1, 7))
    return out


def anchors_standup_f32(anchors_flat: np.ndarray) -> np.ndarray:
    """(A,7) anchors -> the (A,4) float32 boxes the reference passes to bbox_iou (utils.py:213-227, :406-407).

    Parity note: the reference writes the LOW corner into both corner slots (x - l/2, y - w/2 twice for r = 0 anchors,
    x - w/2, y - l/2 twice for r = pi/2), so its anchor boxes are points; that is what is reproduced here."""
    x, y, w, l = anchors_flat[:, 0], anchors_flat[:, 1], anchors_flat[:, 4], anchors_flat[:, 5]
    even = (np.arange(anchors_flat.shape[0]) % 2) == 0
    lo_x = np.where(even, x - l / 2, x - w / 2)
    lo_y = np.where(even, y - w / 2, y - l / 2)
    return np.ascontiguousarray(np.stack([lo_x, lo_y, lo_x, lo_y], axis=1)).astype(np.float32)


def gt_standup_f32(boxes: np.ndarray) -> np.ndarray:
    """(G,7) lidar boxes -> (G,4) float32 axis-aligned bounds of the rotated footprint.

    utils.py:242-255 -> :294-334 (corners: 3x3 yaw rotation times the 3x8 corner template, stored as float32) ->
    :230-239 (min / max over the four bottom corners)."""
    out = np.zeros((boxes.shape[0], 4))
    for i, b in enumerate(boxes):
        w, l, yaw = b[4], b[5], b[6]
        template = np.array([
            [-l / 2, -l / 2, l / 2, l / 2, -l / 2, -l / 2, l / 2, l / 2],
            [w / 2, -w / 2, -w / 2, w / 2, w / 2, -w / 2, -w / 2, w / 2],
            [0, 0, 0, 0, 0.0, 0.0, 0.0, 0.0],
        ])
        rot = np.array([[np.cos(yaw), -np.sin(yaw), 0.0], [np.sin(yaw), np.cos(yaw), 0.0], [0.0, 0.0, 1.0]])
        corners = (np.dot(rot, template) + np.tile(np.array([b[0], b[1], 0.0]), (8, 1)).T).transpose().astype(np.float32)
        foot = corners[0:4, 0:2]
        out[i] = (foot[:, 0].min(), foot[:, 1].min(), foot[:, 0].max(), foot[:, 1].max())
    return np.ascontiguousarray(out).astype(np.float32)


def _device(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("voxelnet_pytorch_b200 runs on CUDA only (sm_100a kernels, no CPU fallback)")
    return torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)


def bbox_iou(box1: np.ndarray, box2: np.ndarray, device=None) -> np.ndarray:
    """(N,4) x (K,4) -> (N,K) float32 overlaps; same values as utils.py:344-373, computed by `vn_bbox_iou`."""
    dev = _device(device)
    a = torch.from_numpy(np.ascontiguousarray(box1, dtype=np.float32)).to(dev)
    b = torch.from_numpy(np.ascontiguousarray(box2, dtype=np.float32)).to(dev)
    return ops.bbox_iou(a, b).cpu().numpy()


class TargetAssigner:
    """Anchors resident on the device; `assign` turns one batch of ground-truth boxes into the three target tensors."""

    def __init__(self, anchors: np.ndarray, cls_name: str = "Car", device=None):
        self.cfg = anchor_config(cls_name)
        self.cls_name = cls_name
        self.device = _device(device)
        flat = np.ascontiguousarray(anchors, dtype=np.float64).reshape(-1, 7)
        self.n_anchors = flat.shape[0]
        self._anchors_host = flat.copy()
        self.anchors = torch.from_numpy(flat).to(self.device)
        self.anchors_standup = torch.from_numpy(anchors_standup_f32(flat)).to(self.device)

    def matches(self, anchors: np.ndarray) -> bool:
        return anchors.size == self._anchors_host.size and np.array_equal(anchors.reshape(-1, 7), self._anchors_host)

    def assign(self, gt_boxes_per_frame):
        """list of (G_f,7) float64 lidar boxes -> (pos, neg) (B,A) f64 and targets (B,A,7) f64 CUDA tensors."""
        counts = [int(b.shape[0]) for b in gt_boxes_per_frame]
        offsets = np.zeros(len(counts) + 1, dtype=np.int32)
        np.cumsum(counts, out=offsets[1:])
        if counts and sum(counts):
            boxes = np.concatenate([np.asarray(b, dtype=np.float64).reshape(-1, 7) for b in gt_boxes_per_frame], axis=0)
        else:
            boxes = np.zeros((0, 7), dtype=np.float64)
        standup = gt_standup_f32(boxes)
        dev = self.device
        return ops.rpn_targets(
            self.anchors_standup, self.anchors, torch.from_numpy(standup).to(dev), torch.from_numpy(boxes).to(dev),
            torch.from_numpy(offsets).to(dev), float(np.float32(self.cfg.rpn_pos_iou)), float(np.float32(self.cfg.rpn_neg_iou)),
            float(self.cfg.anchor_h),
        )


_ASSIGNERS: dict = {}


def generate_targets(labels, feature_map_shape, anchors, cls_name: str = "Car", coordinate: str = "lidar",
                     as_tensors: bool = False, device=None):
    """Drop-in for utils.generate_targets (utils.py:376-462): pos_equal_one, neg_equal_one (B, *fms, 2) and targets
    (B, *fms, 14), float64.  numpy arrays by default like the reference; `as_tensors=True` leaves them on the device
    (model.py:326-331 moves them there anyway)."""
    dev = _device(device)
    key = (cls_name, str(dev))
    assigner = _ASSIGNERS.get(key)
    if assigner is None or not assigner.matches(anchors):
        assigner = _ASSIGNERS[key] = TargetAssigner(anchors, cls_name, dev)
    if assigner.n_anchors != int(np.prod(feature_map_shape)) * 2:
        raise ValueError("anchors (%d) do not match feature_map_shape %s x 2" % (assigner.n_anchors, tuple(feature_map_shape)))
    boxes = label_to_gt_box_3d(labels, cls_name, coordinate)
    if coordinate != "lidar":
        raise NotImplementedError("generate_targets: only coordinate='lidar' (the reference's default and only call)")
    pos, neg, targets = assigner.assign(boxes)
    b = len(boxes)
    pos = pos.view(b, *feature_map_shape, 2)
    neg = neg.view(b, *feature_map_shape, 2)
    targets = targets.view(b, *feature_map_shape, 14)
    if as_tensors:
        return pos, neg, targets
    return pos.cpu().numpy(), neg.cpu().numpy(), targets.cpu().numpy()
