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
