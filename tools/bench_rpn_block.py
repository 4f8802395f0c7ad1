"""GPU box: the stride-1 tail of the reference's RPN block1 / block2 (ConvMD(2,128,128,3,1,1) x 4 / x 5, model.py:212-242) on
the tcgen05 kernel (`ConvMD2d`) against the same layers on stock torch / cuDNN in bf16 channels-last, Car sizes, batch 16.
    python tools/bench_rpn_block.py [out.json]
"""
import json
import sys

sys.path.insert(0, '/root/repo')
import torch
import torch.nn as nn
import torch.nn.functional as F

import voxelnet_pytorch_b200 as vn

dev = torch.device('cuda', 0)
torch.manual_seed(0)


class Ref(nn.Module):
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(128, 128, 3, 1, 1)
        self.batch_norm = nn.BatchNorm2d(128)

    def forward(self, x):
        return F.relu(self.batch_norm(self.conv(x)))


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


out = {}
for name, H, W, L in (("block1_tail", 200, 176, 4), ("block2_tail", 100, 88, 5)):
    B = 16
    ref = nn.Sequential(*[Ref() for _ in range(L)]).to(dev).eval()
    for m in ref.modules():
        if isinstance(m, nn.BatchNorm2d):
            m.running_var.data.uniform_(0.5, 1.5)
            m.running_mean.data.uniform_(-0.2, 0.2)
    x = torch.randn(B, 128, H, W, device=dev).relu().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    ours = vn.RPNBlockTail(ref)
    ref16 = nn.Sequential(*[Ref() for _ in range(L)]).to(dev).eval()
    ref16.load_state_dict(ref.state_dict())
    ref16 = ref16.to(dtype=torch.bfloat16, memory_format=torch.channels_last)
    with torch.no_grad():
        a = ours(x)
        b = ref16(x)
        c = ref(x.float())
        t_ours = timed(lambda: ours(x))
        t_cudnn = timed(lambda: ref16(x))
        t_fp32 = timed(lambda: ref(x.float()))
    flops = L * B * H * W * 9 * 128 * 128 * 2
    err = float((a.float() - c).abs().max() / c.abs().max())
    err16 = float((b.float() - c).abs().max() / c.abs().max())
    out[name] = {"layers": L, "B": B, "H": H, "W": W, "gflop": flops / 1e9, "ms_tcgen05": t_ours, "ms_cudnn_bf16_nhwc": t_cudnn,
                 "ms_torch_fp32": t_fp32, "tflops_tcgen05": flops / t_ours / 1e9, "tflops_cudnn_bf16": flops / t_cudnn / 1e9,
                 "max_rel_err_vs_fp32": err, "cudnn_bf16_max_rel_err_vs_fp32": err16}
    print(name, out[name])
if len(sys.argv) > 1:
    json.dump(out, open(sys.argv[1], 'w'), indent=1)

# ---- the whole RPN half (model.py:212-254, :263-281), Car, batch 16: hybrid vs stock ---------------------------------------
sys.path.insert(0, '/root/repo/tests')
from test_gpu_rpn_conv2d import RefRPN, _randomise  # the reference's RPN architecture restated with its attribute names

g = torch.Generator().manual_seed(1)
ref_rpn = _randomise(RefRPN(), g).to(dev)
hyb = vn.HybridRPN(ref_rpn)
import copy

stock16 = copy.deepcopy(ref_rpn).to(dtype=torch.bfloat16, memory_format=torch.channels_last)
x32 = torch.randn(16, 128, 400, 352, device=dev).relu()
x16 = x32.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
with torch.no_grad():
    p0, r0 = ref_rpn(x32[:2])
    p1, r1 = hyb(x16[:2])
    t_h = timed(lambda: hyb(x16), 10)
    t_s = timed(lambda: stock16(x16), 10)
    t_f = timed(lambda: ref_rpn(x32), 5)
out["rpn_car16"] = {"ms_hybrid_tcgen05": t_h, "ms_stock_bf16_nhwc": t_s, "ms_stock_fp32": t_f,
                    "reg_rel_l2_vs_fp32": float((r1 - r0).norm() / r0.norm()), "prob_max_abs_err": float((p1 - p0).abs().max()),
                    "layers_on_tcgen05": "block1 (5 layers incl. the stride-2 head), block2 (6), deconv1, deconv2, deconv3; block3 too when the batch gives it >= 4 tiles per SM; heads stock"}
print("rpn_car16", out["rpn_car16"])

# ---- points -> RPN maps: the fused front end (bf16 channels-last output) + the hybrid RPN, one step at a time ----------------
from voxelnet_pytorch_b200 import synth

fnet, mnet = vn.FeatureLearningNet('Car').eval().to(dev), vn.MiddleConvNet().eval().to(dev)
fe = vn.VoxelNetFrontEnd(fnet, mnet, 'Car', output_format="nhwc_bf16")
hyb.permute_input_channels(fe.reference_channel_of_output())
pts, offs, max_n = vn.pack_frames(synth.lidar_batch(16, seed0=0))
with torch.no_grad():
    t_all = timed(lambda: hyb(fe.forward_points(pts, offs, max_n)), 10)
    t_fe = timed(lambda: fe.forward_points(pts, offs, max_n), 10)
out["points_to_rpn_maps_car16"] = {"ms_per_batch": t_all, "frames_per_s": 16 / t_all * 1e3, "ms_front_end_alone": t_fe,
                                   "note": "eager launches on one stream; front end returns bf16 channels-last, RPN = HybridRPN"}
print("points_to_rpn_maps_car16", out["points_to_rpn_maps_car16"])
if len(sys.argv) > 1:
    json.dump(out, open(sys.argv[1], 'w'), indent=1)
