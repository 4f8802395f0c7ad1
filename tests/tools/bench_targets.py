"""Timing of the RPN target assignment (Car, 16 frames x 8 boxes): device kernels vs the CPU restatements.

    python tests/tools/bench_targets.py            # prints one JSON object

`device_ms`: vn_rpn_targets alone (CUDA events, inputs resident); `api_ms`: TargetAssigner.assign wall time incl. the
host-side box preparation, uploads and a device sync; `oracle_vectorised_ms`: oracle/targets_np.py (numpy arrays);
`reference_loop_ms_est`: the reference's interpreter loop (oracle bbox_iou_loop) timed on 2,000 anchors and scaled to
70,400 x boxes, IoU part only."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import voxelnet_pytorch_b200  # noqa: E402,F401
from voxelnet_pytorch_b200 import ops, targets as T  # noqa: E402
from oracle import targets_np as OT  # noqa: E402


def main():
    frames, per_frame = 16, 8
    cfg = T.anchor_config("Car")
    rng = np.random.default_rng(0)
    anchors = T.generate_anchors("Car")
    boxes = []
    for _ in range(frames):
        b = np.empty((per_frame, 7))
        b[:, 0], b[:, 1] = rng.uniform(0, 70.4, per_frame), rng.uniform(-40, 40, per_frame)
        b[:, 2], b[:, 3], b[:, 4], b[:, 5] = -1.0, 1.56, 1.6, 3.9
        b[:, 6] = rng.uniform(-1.5, 1.5, per_frame)
        boxes.append(b)
    asg = T.TargetAssigner(anchors, "Car")
    for _ in range(3):
        asg.assign(boxes)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20):
        asg.assign(boxes)
    torch.cuda.synchronize()
    api_ms = (time.perf_counter() - t0) / 20 * 1e3
    dev = asg.device
    allb = np.concatenate(boxes)
    std = torch.from_numpy(T.gt_standup_f32(allb)).to(dev)
    gb = torch.from_numpy(allb).to(dev)
    offs = torch.arange(0, frames * per_frame + 1, per_frame, dtype=torch.int32, device=dev)
    args = (asg.anchors_standup, asg.anchors, std, gb, offs, float(np.float32(cfg.rpn_pos_iou)),
            float(np.float32(cfg.rpn_neg_iou)), cfg.anchor_h)
    for _ in range(3):
        ops.rpn_targets(*args)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(50):
        ops.rpn_targets(*args)
    e1.record()
    torch.cuda.synchronize()
    device_ms = e0.elapsed_time(e1) / 50
    t0 = time.perf_counter()
    OT.generate_targets(boxes, anchors.shape[:2], anchors, cfg.rpn_pos_iou, cfg.rpn_neg_iou, cfg.anchor_h)
    oracle_ms = (time.perf_counter() - t0) * 1e3
    a_std = T.anchors_standup_f32(anchors.reshape(-1, 7))
    g_std = T.gt_standup_f32(boxes[0])
    t0 = time.perf_counter()
    OT.bbox_iou_loop(a_std[:2000], g_std)
    loop_ms = (time.perf_counter() - t0) * 1e3 * (a_std.shape[0] / 2000) * frames
    out_bytes = frames * a_std.shape[0] * (2 + 7) * 8
    print(json.dumps({
        "workload": "Car, %d frames x %d boxes, 70,400 anchors" % (frames, per_frame), "device_ms": device_ms,
        "api_ms": api_ms, "oracle_vectorised_ms": oracle_ms, "reference_loop_ms_est": loop_ms,
        "output_bytes": out_bytes, "device_write_GBps": out_bytes / device_ms / 1e6,
    }))


if __name__ == "__main__":
    main()
