"""Generate tests/golden/targets_car.npz by running the UNMODIFIED reference target assignment (container only; the
generating script of the committed fixture).

    python -m oracle.make_golden_targets

What runs: the reference's `generate_anchors()` (voxelnet/utils.py:104-130), `label_to_gt_box_3d` (:176-209),
`bbox_iou` (:344-373, on a small seeded case and inside generate_targets) and `generate_targets` (:376-462) with its
import-time Car config, on seeded KITTI-format label lines (three frames: several cars and vans with other classes
mixed in, no accepted object at all, one object).  Stored compactly: the label text, the lidar boxes, the index sets of
pos_equal_one / neg_equal_one, the non-zero target rows, a strided sample of the anchors, and the small IoU case.
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_import as R  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
SEED = 77


def synthetic_labels(seed: int = SEED):
    """Three frames of KITTI label_2 lines (type trunc occl alpha bbox*4 h w l x y z ry), camera coordinates."""
    rng = np.random.default_rng(seed)

    def line(kind):
        h, w, l = rng.uniform(1.3, 1.9), rng.uniform(1.4, 1.9), rng.uniform(3.2, 4.6)
        if kind in ("Pedestrian", "Cyclist"):
            h, w, l = rng.uniform(1.5, 1.9), rng.uniform(0.4, 0.8), rng.uniform(0.5, 1.9)
        x, y, z, ry = rng.uniform(-25, 25), rng.uniform(1.2, 2.0), rng.uniform(4, 65), rng.uniform(-np.pi, np.pi)
        return "%s 0.00 0 %.2f 100.00 120.00 300.00 250.00 %.2f %.2f %.2f %.2f %.2f %.2f %.2f\n" % (
            kind, rng.uniform(-np.pi, np.pi), h, w, l, x, y, z, ry)

    frame0 = [line(k) for k in ("Car", "Pedestrian", "Van", "Car", "DontCare", "Car", "Cyclist", "Car", "Van", "Car")]
    frame1 = [line(k) for k in ("Pedestrian", "DontCare", "Tram")]
    frame2 = [line("Car")]
    return [frame0, frame1, frame2]


def main() -> None:
    warnings.filterwarnings("ignore")
    ref = R.load_reference()
    U = ref.utils
    labels = synthetic_labels()
    anchors = U.generate_anchors()
    fms = anchors.shape[:2]
    lab = np.empty(len(labels), dtype=object)
    for i, f in enumerate(labels):
        lab[i] = f
    boxes = U.label_to_gt_box_3d(lab, "Car", "lidar")
    pos, neg, targets = U.generate_targets(lab, fms, anchors)
    rng = np.random.default_rng(SEED + 1)
    lo = rng.uniform(-3, 3, size=(160, 2)).astype(np.float32)
    box1 = np.concatenate([lo, lo + rng.uniform(0, 2.5, size=(160, 2)).astype(np.float32)], axis=1)
    box1[::5, 2:] = box1[::5, :2]  # degenerate boxes like the reference's anchors
    lo = rng.uniform(-3, 3, size=(9, 2)).astype(np.float32)
    box2 = np.concatenate([lo, lo + rng.uniform(0.1, 3, size=(9, 2)).astype(np.float32)], axis=1)
    iou_small = U.bbox_iou(box1, box2)
    t_flat = targets.reshape(len(labels), -1, 7)
    rows = np.argwhere(np.any(t_flat != 0, axis=2))
    out = {
        "seed": np.array(SEED),
        "labels": np.array(["".join(f) for f in labels]),
        "feature_map_shape": np.array(fms),
        "anchors_sample": anchors.reshape(-1, 7)[::97].copy(), "anchors_sum": anchors.sum(axis=(0, 1, 2)),
        "box_counts": np.array([b.shape[0] for b in boxes]), "boxes": np.concatenate(boxes, axis=0),
        "pos_index": np.flatnonzero(pos), "neg_zero_index": np.flatnonzero(neg == 0),
        "pos_sum": np.array(pos.sum()), "neg_sum": np.array(neg.sum()),
        "target_rows": rows, "target_values": t_flat[rows[:, 0], rows[:, 1]],
        "iou_box1": box1, "iou_box2": box2, "iou_small": iou_small,
    }
    path = os.path.join(GOLDEN, "targets_car.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes; positives", int(pos.sum()), "negatives", int(neg.sum()),
          "boxes", out["box_counts"].tolist(), "iou_small range", float(iou_small.min()), float(iou_small.max()))


if __name__ == "__main__":
    main()
