"""Probe: eager launches vs CUDA-graph replay of VoxelNetFrontEnd.forward_points (car16, two alternating batches)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import voxelnet_pytorch_b200 as vn
from voxelnet_pytorch_b200 import synth

dev = torch.device("cuda", 0)
torch.manual_seed(0)
fe = vn.VoxelNetFrontEnd(vn.FeatureLearningNet("Car").eval().to(dev), vn.MiddleConvNet().eval().to(dev), "Car")
sets = [vn.pack_frames(synth.lidar_batch(16, seed0=s)) for s in (0, 500)]
sets.sort(key=lambda p: -p[0].shape[0])
for i in range(6):
    fe.forward_points(*sets[i % 2])
torch.cuda.synchronize()

def timed(fn, n=50):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

print("eager ms/step", timed(lambda i: fe.forward_points(*sets[i % 2])))
graphs = []
side = torch.cuda.Stream()
for k in range(2):
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=side):
        out = fe.forward_points(*sets[k])
    graphs.append(g)
torch.cuda.synchronize()
ref = [None, None]
for k in range(2):
    fe.forward_points(*sets[k]); ref[k] = fe.out.clone()
for i in range(4):
    graphs[i % 2].replay()
torch.cuda.synchronize()
graphs[0].replay(); torch.cuda.synchronize(); print("graph replay equals eager:", torch.equal(fe.out, ref[0]))
print("graph ms/step", timed(lambda i: graphs[i % 2].replay()))
