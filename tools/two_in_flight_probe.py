"""Probe (not part of bench.py): two eval steps in flight on two streams, each with its own front end (own grid,
workspaces, resident outputs), both alternating between the same two batches.  Prints frames/s for one and two
pipelines so that the value of step-level pipelining is known before it is built."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import voxelnet_pytorch_b200 as vn  # noqa: E402
from voxelnet_pytorch_b200 import synth  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    fnet, mnet = vn.FeatureLearningNet("Car").eval().to(dev), vn.MiddleConvNet().eval().to(dev)
    batches = [vn.pack_frames(synth.lidar_batch(16, seed0=s), device=dev) for s in (0, 16)]
    steps = 60

    def run(n_pipes):
        fes = [vn.VoxelNetFrontEnd(fnet, mnet, "Car") for _ in range(n_pipes)]
        streams = [torch.cuda.Stream(device=dev) for _ in range(n_pipes)]
        graphs = [[fe.capture_graph(*b) for b in batches] for fe in fes]
        torch.cuda.synchronize()

        def issue(i):
            p = i % n_pipes
            with torch.cuda.stream(streams[p]):
                graphs[p][(i // n_pipes + p) % 2].replay()

        for i in range(8):
            issue(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in streams:
            s.wait_event(e0)
        for i in range(steps):
            issue(i)
        for s in streams:
            torch.cuda.current_stream().wait_stream(s)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        return {"ms_per_step": ms, "frames_per_s": 16 / ms * 1e3}

    out = {"one_pipeline": run(1), "two_pipelines": run(2), "three_pipelines": run(3)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
