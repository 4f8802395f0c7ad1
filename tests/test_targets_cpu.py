"""CPU: the target-assignment oracle and the product's host-side preparation against the golden vectors frozen from
the unmodified reference (oracle/make_golden_targets.py -> tests/golden/targets_car.npz)."""
import numpy as np
import pytest

from oracle import targets_np as OT
from tests import golden_util as G


def golden():
    return G.load("targets_car.npz")


def golden_boxes(g):
    return np.split(g["boxes"], np.cumsum(g["box_counts"])[:-1])


def golden_labels(g):
    return [s.splitlines(keepends=True) for s in g["labels"]]


def check_targets_against_golden(pos, neg, targets, g, target_rtol=0.0):
    assert pos.dtype == np.float64 and neg.dtype == np.float64 and targets.dtype == np.float64
    fms = tuple(g["feature_map_shape"])
    n_frames = len(g["box_counts"])
    assert pos.shape == (n_frames, *fms, 2) and neg.shape == pos.shape and targets.shape == (n_frames, *fms, 14)
    assert np.array_equal(np.flatnonzero(pos), g["pos_index"]) and pos.sum() == g["pos_sum"]
    assert np.array_equal(np.flatnonzero(neg == 0), g["neg_zero_index"]) and neg.sum() == g["neg_sum"]
    t = targets.reshape(n_frames, -1, 7)
    rows = np.argwhere(np.any(t != 0, axis=2))
    assert np.array_equal(rows, g["target_rows"])
    got = t[rows[:, 0], rows[:, 1]]
    if target_rtol == 0.0:
        assert np.array_equal(got, g["target_values"])
    else:
        assert np.allclose(got, g["target_values"], rtol=target_rtol, atol=0.0)


def test_iou_oracle_matches_reference_golden_bit_exact():
    g = golden()
    want = g["iou_small"]
    assert want.dtype == np.float32 and (want < 0).any() and (want > 1).any()  # the reference's area term at work
    for fn in (OT.bbox_iou, OT.bbox_iou_loop):
        got = fn(g["iou_box1"], g["iou_box2"])
        assert got.dtype == np.float32 and np.array_equal(got.view(np.uint32), want.view(np.uint32)), fn.__name__


def test_generate_targets_oracle_matches_reference_golden():
    from voxelnet_pytorch_b200 import targets as T

    g = golden()
    anchors = T.generate_anchors("Car")
    c = T.anchor_config("Car")
    pos, neg, targets = OT.generate_targets(golden_boxes(g), tuple(g["feature_map_shape"]), anchors,
                                            c.rpn_pos_iou, c.rpn_neg_iou, c.anchor_h)
    check_targets_against_golden(pos, neg, targets, g)


def test_host_preparation_matches_reference_golden():
    """Anchors, label parsing + camera->lidar boxes and both stand-up box builders: the values the device receives."""
    from voxelnet_pytorch_b200 import targets as T

    g = golden()
    anchors = T.generate_anchors("Car")
    assert anchors.shape == (200, 176, 2, 7) and anchors.dtype == np.float64
    assert np.array_equal(anchors.reshape(-1, 7)[::97], g["anchors_sample"])
    assert np.array_equal(anchors.sum(axis=(0, 1, 2)), g["anchors_sum"])
    boxes = T.label_to_gt_box_3d(golden_labels(g), "Car", "lidar")
    assert [b.shape[0] for b in boxes] == g["box_counts"].tolist()
    assert np.array_equal(np.concatenate(boxes, axis=0), g["boxes"])
    flat = anchors.reshape(-1, 7)
    assert np.array_equal(T.anchors_standup_f32(flat), OT.anchors_standup(flat).astype(np.float32))
    assert np.array_equal(T.gt_standup_f32(g["boxes"]), OT.gt_standup(g["boxes"]).astype(np.float32))
    assert T.gt_standup_f32(np.zeros((0, 7))).shape == (0, 4)
    ped = T.generate_anchors("Pedestrian")
    assert ped.shape == (100, 120, 2, 7) and ped[0, 0, 1, 6] == 90 / 180 * np.pi and ped[3, 5, 0, 5] == 0.8
    with pytest.raises(ValueError):
        T.anchor_config("Tram")


def test_label_classes_and_unknown_class_accepts_everything():
    from voxelnet_pytorch_b200 import targets as T

    labels = golden_labels(golden())
    assert [b.shape[0] for b in T.label_to_gt_box_3d(labels, "Pedestrian", "lidar")] == [1, 1, 0]
    assert [b.shape[0] for b in T.label_to_gt_box_3d(labels, "Cyclist", "lidar")] == [1, 0, 0]
    assert [b.shape[0] for b in T.label_to_gt_box_3d(labels, "anything", "camera")] == [10, 3, 1]
    cam = T.label_to_gt_box_3d(labels, "Car", "camera")[2][0]
    assert cam[3:6].tolist() == [float(v) for v in labels[2][0].split()[8:11]]  # h, w, l untouched


def test_product_targets_fail_loudly_without_cuda():
    import torch
    from voxelnet_pytorch_b200 import targets as T

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError):
        T.bbox_iou(np.zeros((2, 4), np.float32), np.zeros((1, 4), np.float32))
    with pytest.raises(RuntimeError):
        T.generate_targets([[]], (200, 176), T.generate_anchors("Car"))


@pytest.mark.parametrize("seed", range(4))
def test_iou_array_restatement_equals_the_literal_loop(seed):
    """The array form of the IoU (what the GPU parity tests compare against at full size) has the bits of the literal
    double loop on random, degenerate, touching and disjoint boxes."""
    rng = np.random.default_rng(seed)
    lo = rng.uniform(-5, 5, size=(60, 2)).astype(np.float32)
    box1 = np.concatenate([lo, lo + rng.uniform(0, 4, size=(60, 2)).astype(np.float32)], axis=1)
    box1[::3, 2:] = box1[::3, :2]                      # points, like the reference's anchors
    lo = rng.uniform(-5, 5, size=(7, 2)).astype(np.float32)
    box2 = np.concatenate([lo, lo + rng.uniform(0, 4, size=(7, 2)).astype(np.float32)], axis=1)
    box2[0] = box1[1]                                  # identical boxes
    box2[1] = box1[2] + np.float32(100)                # far away: iw <= 0
    box1[4] = (0, 1, 0, 1)                             # anchor-area term (y1 - x1 + 1) = 2, (y2 - y1 + 1) = 1
    a, b = OT.bbox_iou(box1, box2), OT.bbox_iou_loop(box1, box2)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert (a[:, 1] == 0).all()


def test_generate_targets_oracle_edge_cases():
    """No boxes at all -> everything negative; one box far outside the range -> its arg-max anchor still becomes
    positive only if its IoU is > 0."""
    from voxelnet_pytorch_b200 import targets as T

    anchors = T.generate_anchors("Pedestrian")
    c = T.anchor_config("Pedestrian")
    fms = anchors.shape[:2]
    pos, neg, tg = OT.generate_targets([np.zeros((0, 7))], fms, anchors, c.rpn_pos_iou, c.rpn_neg_iou, c.anchor_h)
    assert pos.sum() == 0 and neg.sum() == neg.size and not tg.any()
    far = np.array([[500.0, 500.0, -1.0, 1.7, 0.6, 0.8, 0.3]])
    pos, neg, tg = OT.generate_targets([far], fms, anchors, c.rpn_pos_iou, c.rpn_neg_iou, c.anchor_h)
    assert pos.sum() == 0 and neg.sum() == neg.size
