"""TEST INFRASTRUCTURE - CPU restatement of the reference's RPN target assignment (never imported by the product).

Follows voxelnet/utils.py: `bbox_iou` (:344-373), `anchor_to_standup_box2d` (:213-227), `center_to_corner_box_2d` /
`center_to_corner_box_3d` / `corner_to_standup_box2d` (:242-255, :294-334, :230-239) and the per-frame body of
`generate_targets` (:395-462).  The interpreter double loop of `bbox_iou` is restated as float32 array expressions in
the same operation order (every numpy ufunc rounds once, so the values are the loop's values); `bbox_iou_loop` keeps
the literal loop for small cases.

Pinned against the live reference: `oracle/make_golden_targets.py` runs the unmodified `generate_anchors`,
`label_to_gt_box_3d`, `bbox_iou` and `generate_targets` on seeded KITTI-format label lines and stores their outputs in
`tests/golden/targets_car.npz`; `tests/test_oracle_golden.py` checks this file against that fixture and
`tests/test_oracle_vs_reference.py` against the reference itself when it is present.
"""
from __future__ import annotations

import numpy as np


def bbox_iou_loop(box1: np.ndarray, box2: np.ndarray) -> np.ndarray:
    """utils.py:344-373 as written (float32 scalars, one entry at a time).  Small inputs only."""
    n_rows, n_cols = box1.shape[0], box2.shape[0]
    out = np.zeros((n_rows, n_cols), dtype=np.float32)
    one = np.float32(1)
    for k in range(n_cols):
        g = box2[k]
        g_area = (g[2] - g[0] + one) * (g[3] - g[1] + one)
        for n in range(n_rows):
            a = box1[n]
            iw = min(a[2], g[2]) - max(a[0], g[0]) + one
            if iw > 0:
                ih = min(a[3], g[3]) - max(a[1], g[1]) + one
                if ih > 0:
                    # the anchor-area term as the reference has it (:366-367): (y1 - x1 + 1) * (y2 - y1 + 1)
                    ua = (a[1] - a[0] + one) * (a[3] - a[1] + one) + g_area - iw * ih
                    with np.errstate(divide="ignore"):
                        out[n, k] = iw * ih / ua
    return out


def bbox_iou(box1: np.ndarray, box2: np.ndarray) -> np.ndarray:
    """The same values, one float32 array expression per operation of the loop."""
    a = np.ascontiguousarray(box1, dtype=np.float32)[:, None, :]
    g = np.ascontiguousarray(box2, dtype=np.float32)[None, :, :]
    one = np.float32(1)
    g_area = (g[..., 2] - g[..., 0] + one) * (g[..., 3] - g[..., 1] + one)
    iw = np.minimum(a[..., 2], g[..., 2]) - np.maximum(a[..., 0], g[..., 0]) + one
    ih = np.minimum(a[..., 3], g[..., 3]) - np.maximum(a[..., 1], g[..., 1]) + one
    inter = iw * ih
    ua = (a[..., 1] - a[..., 0] + one) * (a[..., 3] - a[..., 1] + one) + g_area - inter
    with np.errstate(divide="ignore", invalid="ignore"):
        v = inter / ua
    return np.where((iw > 0) & (ih > 0), v, np.float32(0)).astype(np.float32)


def anchors_standup(anchors_flat: np.ndarray) -> np.ndarray:
    """utils.py:213-227 on columns (x, y, w, l) of the (A,7) anchors: both corner slots receive the low corner."""
    sel = anchors_flat[:, [0, 1, 4, 5]]
    out = np.zeros_like(sel)
    for start, (dx, dy) in ((0, (3, 2)), (1, (2, 3))):
        rows = slice(start, None, 2)
        out[rows, 0] = out[rows, 2] = sel[rows, 0] - sel[rows, dx] / 2
        out[rows, 1] = out[rows, 3] = sel[rows, 1] - sel[rows, dy] / 2
    return out


def gt_standup(boxes: np.ndarray) -> np.ndarray:
    """(G,7) lidar boxes -> (G,4) float64 array holding float32 values: bounds of the four bottom corners."""
    corners = np.zeros((boxes.shape[0], 8, 3), dtype=np.float32)
    for i in range(boxes.shape[0]):
        x, y, w, l, yaw = boxes[i, 0], boxes[i, 1], boxes[i, 4], boxes[i, 5], boxes[i, 6]
        h = 0.0  # center_to_corner_box_2d zeroes z and h (:250-251)
        shape = np.array([
            [-l / 2, -l / 2, l / 2, l / 2, -l / 2, -l / 2, l / 2, l / 2],
            [w / 2, -w / 2, -w / 2, w / 2, w / 2, -w / 2, -w / 2, w / 2],
            [0, 0, 0, 0, h, h, h, h],
        ])
        rot = np.array([[np.cos(yaw), -np.sin(yaw), 0.0], [np.sin(yaw), np.cos(yaw), 0.0], [0.0, 0.0, 1.0]])
        corners[i] = (np.dot(rot, shape) + np.tile(np.array([x, y, 0.0]), (8, 1)).T).transpose()
    foot = corners[:, 0:4, 0:2]
    out = np.zeros((boxes.shape[0], 4))
    out[:, 0] = np.min(foot[:, :, 0], axis=1)
    out[:, 1] = np.min(foot[:, :, 1], axis=1)
    out[:, 2] = np.max(foot[:, :, 0], axis=1)
    out[:, 3] = np.max(foot[:, :, 1], axis=1)
    return out


def generate_targets(gt_boxes_per_frame, feature_map_shape, anchors, pos_iou: float, neg_iou: float, anchor_h: float,
                     iou_fn=bbox_iou):
    """utils.py:376-462 from the per-frame (G,7) lidar boxes on (label parsing is host-side text work)."""
    batch = len(gt_boxes_per_frame)
    flat = anchors.reshape(-1, 7)
    diag = np.sqrt(flat[:, 4] ** 2 + flat[:, 5] ** 2)
    pos = np.zeros((batch, *feature_map_shape, 2))
    neg = np.zeros((batch, *feature_map_shape, 2))
    targets = np.zeros((batch, *feature_map_shape, 14))
    a_std = np.ascontiguousarray(anchors_standup(flat)).astype(np.float32)
    for b, gt in enumerate(gt_boxes_per_frame):
        iou = iou_fn(a_std, np.ascontiguousarray(gt_standup(gt)).astype(np.float32))
        best_anchor = np.argmax(iou.T, axis=1)
        best_box = np.arange(iou.shape[1])
        keep = iou.T[best_box, best_anchor] > 0
        best_anchor, best_box = best_anchor[keep], best_box[keep]
        over_a, over_g = np.where(iou > pos_iou)
        neg_a = np.where(np.sum(iou < neg_iou, axis=1) == iou.shape[1])[0]
        pos_a, first = np.unique(np.concatenate([over_a, best_anchor]), return_index=True)
        pos_g = np.concatenate([over_g, best_box])[first]
        iy, ix, ir = np.unravel_index(pos_a, (*feature_map_shape, 2))
        pos[b, iy, ix, ir] = 1
        g, an = gt[pos_g], flat[pos_a]
        targets[b, iy, ix, ir * 7 + 0] = (g[:, 0] - an[:, 0]) / diag[pos_a]
        targets[b, iy, ix, ir * 7 + 1] = (g[:, 1] - an[:, 1]) / diag[pos_a]
        targets[b, iy, ix, ir * 7 + 2] = (g[:, 2] - an[:, 2]) / anchor_h
        for j in (3, 4, 5):
            targets[b, iy, ix, ir * 7 + j] = np.log(g[:, j] / an[:, j])
        targets[b, iy, ix, ir * 7 + 6] = g[:, 6] - an[:, 6]
        iy, ix, ir = np.unravel_index(neg_a, (*feature_map_shape, 2))
        neg[b, iy, ix, ir] = 1
    return pos, neg, targets
