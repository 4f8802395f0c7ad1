"""GPU parity: vn_bbox_iou / vn_rpn_targets (through the C ABI) against the oracle and the reference golden vectors.

Bars: IoU bit-exact (float32); positive / negative index sets identical; regression targets float64 within 4 ulp
(rtol 1e-15: every term is IEEE-exact except log(), 1 ulp on the device and in numpy)."""
import numpy as np
import pytest
import torch

from oracle import targets_np as OT
from tests.test_targets_cpu import check_targets_against_golden, golden, golden_boxes, golden_labels

pytestmark = pytest.mark.gpu
TARGET_RTOL = 1e-15


def random_boxes(rng, n, cfg):
    """Lidar boxes spread over the class's range (x, y, z, h, w, l, r)."""
    b = np.empty((n, 7))
    b[:, 0] = rng.uniform(cfg.x_range[0], cfg.x_range[1], n)
    b[:, 1] = rng.uniform(cfg.y_range[0], cfg.y_range[1], n)
    b[:, 2] = rng.uniform(-2.0, -0.5, n)
    b[:, 3] = cfg.anchor_h * rng.uniform(0.8, 1.2, n)
    b[:, 4] = cfg.anchor_w * rng.uniform(0.8, 1.2, n)
    b[:, 5] = cfg.anchor_l * rng.uniform(0.8, 1.2, n)
    b[:, 6] = rng.uniform(-np.pi / 2, np.pi / 2, n)
    return b


def test_bbox_iou_bit_exact(cuda_device):
    from voxelnet_pytorch_b200 import targets as T

    g = golden()
    got = T.bbox_iou(g["iou_box1"], g["iou_box2"])
    assert got.dtype == np.float32 and np.array_equal(got.view(np.uint32), g["iou_small"].view(np.uint32))
    # full size: all 70,400 Car anchors (as the reference builds them) against 13 boxes
    flat = T.generate_anchors("Car").reshape(-1, 7)
    a_std = T.anchors_standup_f32(flat)
    g_std = T.gt_standup_f32(random_boxes(np.random.default_rng(3), 13, T.anchor_config("Car")))
    got, want = T.bbox_iou(a_std, g_std), OT.bbox_iou(a_std, g_std)
    assert got.shape == (70400, 13) and np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.isinf(want).any() or (want > 0.6).any()
    assert T.bbox_iou(a_std[:0], g_std).shape == (0, 13) and T.bbox_iou(a_std[:5], g_std[:0]).shape == (5, 0)


def test_generate_targets_matches_reference_golden(cuda_device):
    from voxelnet_pytorch_b200 import targets as T

    g = golden()
    anchors = T.generate_anchors("Car")
    pos, neg, targets = T.generate_targets(np.array(golden_labels(g), dtype=object), (200, 176), anchors)
    check_targets_against_golden(pos, neg, targets, g, target_rtol=TARGET_RTOL)
    p2, n2, t2 = T.generate_targets(golden_labels(g), (200, 176), anchors, as_tensors=True)
    assert p2.is_cuda and t2.shape == (3, 200, 176, 14) and t2.dtype == torch.float64
    assert np.array_equal(p2.cpu().numpy(), pos) and np.array_equal(n2.cpu().numpy(), neg)
    assert np.array_equal(t2.cpu().numpy(), targets)  # deterministic
    with pytest.raises(ValueError):
        T.generate_targets(golden_labels(g), (100, 176), anchors)


@pytest.mark.parametrize("cls_name,counts", [("Car", [9, 0, 1, 30]), ("Pedestrian", [0, 17, 3]), ("Cyclist", [1100, 0, 2])])
def test_rpn_targets_match_oracle(cuda_device, cls_name, counts):
    """Random boxes incl. empty frames and a frame with more boxes than one shared-memory pass holds (1100 > 1024)."""
    from voxelnet_pytorch_b200 import targets as T

    cfg = T.anchor_config(cls_name)
    rng = np.random.default_rng(len(counts) * 100 + counts[0])
    anchors = T.generate_anchors(cls_name)
    fms = anchors.shape[:2]
    boxes = [random_boxes(rng, n, cfg) for n in counts]
    if counts[-1] >= 2:  # two identical boxes: ties between boxes resolve to the lower-numbered one
        boxes[-1][1] = boxes[-1][0]
    pos, neg, targets = T.TargetAssigner(anchors, cls_name).assign(boxes)
    pos, neg, targets = pos.cpu().numpy(), neg.cpu().numpy(), targets.cpu().numpy()
    wp, wn, wt = OT.generate_targets(boxes, fms, anchors, np.float32(cfg.rpn_pos_iou), np.float32(cfg.rpn_neg_iou), cfg.anchor_h)
    assert np.array_equal(pos.reshape(wp.shape), wp) and np.array_equal(neg.reshape(wn.shape), wn)
    assert wp.sum() > 0 and wn.sum() < wn.size
    assert np.allclose(targets.reshape(wt.shape), wt, rtol=TARGET_RTOL, atol=0.0)
    for b, n in enumerate(counts):
        if n == 0:
            assert pos[b].sum() == 0 and neg[b].sum() == neg[b].size and not targets[b].any()
