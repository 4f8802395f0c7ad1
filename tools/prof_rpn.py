"""GPU box: where do the 3.7 ms of HybridRPN (Car, batch 16) go?  CUDA events around its sections."""
import sys

sys.path.insert(0, '/root/repo')
sys.path.insert(0, '/root/repo/tests')
import torch

import voxelnet_pytorch_b200 as vn
from test_gpu_rpn_conv2d import RefRPN, _randomise

dev = torch.device('cuda', 0)
g = torch.Generator().manual_seed(1)
rpn = vn.HybridRPN(_randomise(RefRPN(), g).to(dev))
x = torch.randn(16, 128, 400, 352, device=dev).relu().to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
marks = []


def ev(name):
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    marks.append((name, e))


def run():
    marks.clear()
    plans = rpn.__dict__.setdefault("_plans", vn.rpn._PlanCache())
    ev("start")
    y = rpn.block1(x, plans); ev("block1 (5 layers, tcgen05)")
    B, _, H, W = y.shape
    cat = torch.empty((B, H, W, 768), dtype=torch.bfloat16, device=dev); ev("alloc cat")
    rpn.deconv_fast[0].forward_nhwc(y.permute(0, 2, 3, 1), out=cat, plans=plans, channel_offset=512); ev("deconv1 (tcgen05, into cat)")
    y = rpn.block2(y, plans); ev("block2 (6 layers, tcgen05)")
    d2 = rpn.deconv_stock[1](y); ev("deconv2 (stock)")
    cat[..., 256:512] = d2.permute(0, 2, 3, 1); ev("copy deconv2 into cat")
    y = rpn.block3(y, plans); ev("block3 (6 layers, stock at this batch)")
    d3 = rpn.deconv_stock[2](y); ev("deconv3 (stock)")
    cat[..., :256] = d3.permute(0, 2, 3, 1); ev("copy deconv3 into cat")
    h = rpn.heads(cat.permute(0, 3, 1, 2)); ev("heads 1x1 (stock, fused)")
    o = h.float(); p = torch.sigmoid(o[:, :2]).contiguous(); r = o[:, 2:].contiguous(); ev("float / sigmoid / split")


with torch.no_grad():
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    acc = {}
    for _ in range(10):
        run()
        torch.cuda.synchronize()
        for (n0, e0), (n1, e1) in zip(marks[:-1], marks[1:]):
            acc[n1] = acc.get(n1, 0.0) + e0.elapsed_time(e1) / 10
for k, v in acc.items():
    print("%-45s %.3f ms" % (k, v))
print("total %.3f ms" % sum(acc.values()))
