"""Same-box GPU comparator: the reference's module graph for the hot path built from STOCK torch.nn
modules (cuBLAS / cuDNN / torch sparse kernels), timed on the same B200 and the same batch as `bench.py`'s car16.

    python tools/gpu_comparator.py [--frames 16] [--steps 10]

What it runs (architecture of voxelnet/model.py:60-108, 111-167, 206-262, restated with nn.Linear / nn.BatchNorm1d /
torch.sparse_coo_tensor(...).to_dense() / nn.Conv3d / nn.BatchNorm3d; eval mode): VFE-1 -> VFE-2 -> voxel-wise max ->
sparse-to-dense -> 3 x (Conv3d + BN + ReLU) -> reshape, on voxel features already on the device (the reference
voxelizes with numpy on the host; that part is NOT in this number, so the comparison favours the comparator).
Modes: fp32 with TF32 convolutions (torch's default), and bf16 autocast.  This is library code running the reference's
graph - "the existing Blackwell path" - not this package's kernels; the package's own step is timed beside it through
`VoxelNetFrontEnd.forward_points` (which does include the voxelization).  Prints one JSON object.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import voxelnet_pytorch_b200 as vn  # noqa: E402
from voxelnet_pytorch_b200 import synth  # noqa: E402


class StockVFE(nn.Module):
    def __init__(self, cin, cout):
        super().__init__()
        self.units = cout // 2
        self.fcn = nn.Sequential(nn.Linear(cin, self.units), nn.ReLU())
        self.bn = nn.BatchNorm1d(self.units)

    def forward(self, x, mask):
        k, t, _ = x.shape
        pw = self.bn(self.fcn(x).reshape(k * t, self.units)).reshape(k, t, self.units)
        agg = pw.max(dim=1, keepdim=True).values.expand(k, t, self.units)
        return torch.cat([pw, agg], dim=2) * mask


class StockFeatureNet(nn.Module):
    def __init__(self, grid):
        super().__init__()
        self.grid = grid
        self.vfe_1, self.vfe_2 = StockVFE(7, 32), StockVFE(32, 128)

    def forward(self, feature, coordinate, batch):
        mask = (feature.max(dim=2, keepdim=True).values != 0).to(feature.dtype)
        x = self.vfe_2(self.vfe_1(feature, mask), mask)
        vw = x.max(dim=1).values
        g = self.grid
        size = (batch, g.depth, g.height, g.width, 128)
        return torch.sparse_coo_tensor(coordinate.t(), vw.float(), size).to_dense()


class StockConv(nn.Module):
    def __init__(self, cin, cout, stride, padding):
        super().__init__()
        self.conv = nn.Conv3d(cin, cout, 3, stride, padding)
        self.batch_norm = nn.BatchNorm3d(cout)

    def forward(self, x):
        return torch.relu(self.batch_norm(self.conv(x)))


class StockMiddle(nn.Module):
    def __init__(self):
        super().__init__()
        self.middle_layer = nn.Sequential(*[StockConv(c["cin"], c["cout"], c["stride"], c["padding"])
                                            for c in vn.config.MIDDLE_LAYERS])

    def forward(self, dense):
        b, _, h, w, _ = dense.shape
        return self.middle_layer(dense.permute(0, 4, 1, 2, 3)).reshape(b, -1, h, w)


def timed(fn, steps, warmup=3):
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=16)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--target", default="Car")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    grid = vn.grid_config(args.target)
    fnet, mnet = vn.FeatureLearningNet(args.target).eval(), vn.MiddleConvNet().eval()
    g = torch.Generator().manual_seed(1)
    for m in list(fnet.modules()) + list(mnet.modules()):
        if isinstance(m, nn.modules.batchnorm._BatchNorm):
            m.weight.data = 0.5 + torch.rand(m.weight.shape, generator=g)
            m.bias.data = (torch.rand(m.bias.shape, generator=g) - 0.5) * 0.6
            m.running_mean.data = (torch.rand(m.running_mean.shape, generator=g) - 0.5) * 0.4
            m.running_var.data = 0.5 + torch.rand(m.running_var.shape, generator=g)
    sf, sm = StockFeatureNet(grid).eval(), StockMiddle().eval()
    sf.load_state_dict(fnet.state_dict())
    sm.load_state_dict(mnet.state_dict())
    sf, sm = sf.to(dev), sm.to(dev)
    fe = vn.VoxelNetFrontEnd(fnet.to(dev), mnet.to(dev), args.target)

    frames = synth.lidar_batch(args.frames, seed0=0)
    points, offsets, max_n = vn.pack_frames(frames, device=dev)
    vb = vn.voxelize_batch(points, offsets, max_n, args.target)
    k = int(vb.voxel_offsets[-1])
    feature, coordinate = vb.feature[:k].contiguous(), vb.coordinate[:k].contiguous()

    out = {"workload": "%s, %d frames, %d voxels (voxel features resident; stock path excludes voxelization)"
                       % (args.target, args.frames, k), "torch": torch.__version__,
           "cudnn": torch.backends.cudnn.version(), "tf32_conv": torch.backends.cudnn.allow_tf32}
    with torch.no_grad():
        mine = fe.forward_points(points, offsets, max_n).float().clone()
        want = sm(sf(feature, coordinate, args.frames))
        out["max_abs_diff_vs_stock_over_max"] = float((mine - want).abs().max() / want.abs().max())
        del want, mine

        def stock_fp32():
            return sm(sf(feature, coordinate, args.frames))

        def stock_bf16():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return sm(sf(feature, coordinate, args.frames))

        def stock_bf16_channels_last():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                dense = sf(feature, coordinate, args.frames)
                b, _, h, w, _ = dense.shape
                x = dense.permute(0, 4, 1, 2, 3)  # already channels_last_3d in memory
                return sm.middle_layer(x).reshape(b, -1, h, w)

        for name, fn in (("stock_fp32_tf32conv", stock_fp32), ("stock_bf16_autocast", stock_bf16),
                         ("stock_bf16_autocast_channels_last", stock_bf16_channels_last)):
            try:
                ms = timed(fn, args.steps)
                out[name] = {"ms_per_step": ms, "frames_per_s": args.frames / ms * 1e3}
            except RuntimeError as exc:  # e.g. out of memory in a cuDNN workspace
                out[name] = {"error": str(exc)[:160]}
            torch.cuda.empty_cache()
        ms = timed(lambda: fe.forward_points(points, offsets, max_n), max(args.steps, 30))
        out["this_package_incl_voxelization"] = {"ms_per_step": ms, "frames_per_s": args.frames / ms * 1e3}
        out["peak_memory_GB"] = torch.cuda.max_memory_allocated() / 1e9
    print(json.dumps(out))


if __name__ == "__main__":
    main()
