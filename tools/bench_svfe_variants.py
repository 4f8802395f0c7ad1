"""A/B harness for changes to the voxel -> SVFE kernel: prints the svfe / voxelize stage times (eager, CUDA events between the
stages), the one-step-at-a-time graph-replay time and a checksum of the output (bit-identical variants print the same one).
While variants were compared the library read VN_SVFE_VARIANT once per process (history); the shipped
build has one variant and ignores it.  Usage: python tools/bench_svfe_variants.py [reps]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import voxelnet_pytorch_b200 as vn  # noqa: E402
from voxelnet_pytorch_b200 import synth  # noqa: E402
from voxelnet_pytorch_b200.pipeline import StageTimer  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
dev = torch.device("cuda", 0)
torch.manual_seed(0)
fnet, mnet = vn.FeatureLearningNet("Car").eval().to(dev), vn.MiddleConvNet().eval().to(dev)
fe = vn.VoxelNetFrontEnd(fnet, mnet, "Car")
batches = []
for s in (0, 500):
    fr = synth.lidar_batch(16, seed0=s)
    offs = np.zeros(17, dtype=np.int32)
    np.cumsum([f.shape[0] for f in fr], out=offs[1:])
    batches.append((torch.from_numpy(np.concatenate(fr, 0)).to(dev), torch.from_numpy(offs).to(dev), max(f.shape[0] for f in fr)))
batches.sort(key=lambda b: -b[0].shape[0])
for i in range(6):
    out = fe.forward_points(*batches[i % 2])
torch.cuda.synchronize()
check = float(out.double().abs().sum())
t = StageTimer()
for i in range(reps):
    fe.forward_points(*batches[i % 2], timer=t)
torch.cuda.synchronize()
graphs = [fe.capture_graph(*b) for b in batches]
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for i in range(6):
    graphs[i % 2].replay()
torch.cuda.synchronize()
e0.record()
for i in range(reps * 2):
    graphs[i % 2].replay()
e1.record()
torch.cuda.synchronize()
print(json.dumps({"variant": os.environ.get("VN_SVFE_VARIANT", "0"), "f32_store": os.environ.get("VN_CONV_F32_STORE", "0"),
                  "stages_ms": {k: round(v[0], 4) for k, v in t.summary().items()}, "svfe_ms": t.summary()["svfe"][0], "voxelize_ms": t.summary()["voxelize"][0],
                  "step_ms": e0.elapsed_time(e1) / (reps * 2), "checksum": check}))
