"""GPU: vn_rpn_targets_static (the allocation-free variant of vn_rpn_targets, scratch in the library's static device
array) against vn_rpn_targets (caller workspace), through the C ABI, on two calls in a row.

Bar: bit-identical outputs (the same four kernels run on both paths; the tests of vn_rpn_targets itself hold it to the
oracle: index sets identical, targets within rtol 1e-15 - tests/test_gpu_targets.py).  The second call uses fewer frames
and fewer boxes than the first, so the array still holds the first call's contents where the second one looks.

Collected last on purpose; passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(cls_name, counts, seed, dev):
    from tests.test_gpu_targets import random_boxes
    from voxelnet_pytorch_b200 import targets as T

    cfg = T.anchor_config(cls_name)
    rng = np.random.default_rng(seed)
    flat = np.ascontiguousarray(T.generate_anchors(cls_name), dtype=np.float64).reshape(-1, 7)
    boxes = [random_boxes(rng, n, cfg) for n in counts]
    offsets = np.zeros(len(counts) + 1, dtype=np.int32)
    np.cumsum(counts, out=offsets[1:])
    allb = np.concatenate(boxes, 0) if sum(counts) else np.zeros((0, 7))
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)  # noqa: E731
    return dict(
        anchors_standup=t(T.anchors_standup_f32(flat)), anchors=t(flat), gt_standup=t(T.gt_standup_f32(allb)), gt_boxes=t(allb),
        gt_offsets=t(offsets), pos_iou=float(np.float32(cfg.rpn_pos_iou)), neg_iou=float(np.float32(cfg.rpn_neg_iou)),
        anchor_h=float(cfg.anchor_h),
    )


def _outputs(B, A, dev, fill):
    return (torch.full((B, A), fill, dtype=torch.float64, device=dev), torch.full((B, A), fill, dtype=torch.float64, device=dev),
            torch.full((B, A, 7), fill, dtype=torch.float64, device=dev))


def test_static_scratch_entry_point_matches_the_workspace_entry_point_twice(cuda_device):
    from voxelnet_pytorch_b200 import _cabi, ops

    dev = cuda_device
    assert ops.rpn_targets_static_capacity_bytes() == 16 << 20
    calls = [("Car", [9, 0, 1, 30, 1100, 2], 11), ("Car", [3, 5], 21)]  # 2nd: fewer frames and boxes -> stale scratch beyond its extent
    for n_call, (cls_name, counts, seed) in enumerate(calls):
        kw = _inputs(cls_name, counts, seed, dev)
        B, A = len(counts), kw["anchors"].shape[0]
        want = ops.rpn_targets(**kw)
        outs = _outputs(B, A, dev, float("nan"))  # every element must be written
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(dev)[0]
        launches0 = _cabi.launch_count()
        got = ops.rpn_targets_static(**kw, pos=outs[0], neg=outs[1], targets=outs[2])
        torch.cuda.synchronize()
        assert _cabi.launch_count() - launches0 == 4
        if n_call > 0:  # (the first call may load the module's kernels / resolve the symbol lazily)
            assert torch.cuda.mem_get_info(dev)[0] >= free0, "device memory was taken during the call"
        for g, w, name in zip(got, want, ("pos_equal_one", "neg_equal_one", "targets")):
            assert torch.equal(g.view(torch.int64), w.view(torch.int64)), "call %d: %s differs" % (n_call, name)
        assert float(got[0].sum()) > 0 and float(got[1].sum()) < got[1].numel()

    # a frame without boxes after calls that had some, on a side stream: all negatives, zero targets
    kw = _inputs("Pedestrian", [0, 4], 13, dev)
    B, A = 2, kw["anchors"].shape[0]
    want = ops.rpn_targets(**kw)
    outs = _outputs(B, A, dev, -7.0)
    torch.cuda.synchronize()
    side = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(side):
        got = ops.rpn_targets_static(**kw, pos=outs[0], neg=outs[1], targets=outs[2])
    side.synchronize()
    assert all(torch.equal(g, w) for g, w in zip(got, want))
    assert float(got[0][0].sum()) == 0 and float(got[1][0].sum()) == A and not bool(got[2][0].any())


def test_static_scratch_entry_point_rejects_what_does_not_fit(cuda_device):
    from voxelnet_pytorch_b200 import _cabi, ops

    dev = cuda_device
    lib = _cabi.lib()
    kw = _inputs("Car", [1] * 30, 5, dev)  # 30 x 70,400 anchors x 8 bytes > 16 MiB
    B, A = 30, kw["anchors"].shape[0]
    assert int(lib.vn_rpn_targets_workspace_bytes(B, A, 30)) > ops.rpn_targets_static_capacity_bytes()
    outs = _outputs(B, A, dev, 3.0)
    launches0 = _cabi.launch_count()
    with pytest.raises(_cabi.VnError, match="static array holds"):
        ops.rpn_targets_static(**kw, pos=outs[0], neg=outs[1], targets=outs[2])
    rc = lib.vn_rpn_targets_static(
        C.c_void_p(kw["anchors_standup"].data_ptr()), C.c_void_p(kw["anchors"].data_ptr()), A, C.c_void_p(kw["gt_standup"].data_ptr()),
        C.c_void_p(kw["gt_boxes"].data_ptr()), C.c_void_p(kw["gt_offsets"].data_ptr()), B, 30, kw["pos_iou"], kw["neg_iou"], kw["anchor_h"],
        C.c_void_p(outs[0].data_ptr()), C.c_void_p(outs[1].data_ptr()), C.c_void_p(outs[2].data_ptr()), None)
    assert rc == -2  # VN_ERR_WORKSPACE
    torch.cuda.synchronize()
    assert _cabi.launch_count() == launches0 and float(outs[0].min()) == 3.0  # nothing was launched, nothing written
    # the largest Car batch that fits (29 frames) runs
    kw = _inputs("Car", [1] * 29, 6, dev)
    outs = _outputs(29, A, dev, float("nan"))
    got = ops.rpn_targets_static(**kw, pos=outs[0], neg=outs[1], targets=outs[2])
    want = ops.rpn_targets(**kw)
    assert all(torch.equal(g, w) for g, w in zip(got, want))
