"""Print the observed gradient-parity numbers of the training tests (how far each is from its stated bar)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import voxelnet_pytorch_b200 as vn
from tests.helpers import load_state
from tests.test_oracle_golden import _train_golden_inputs
from voxelnet_pytorch_b200.utils import pack_frames

dev = torch.device("cuda", 0)
gd = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "golden")
for rep in range(3):
    g, points, target, sd, dout = _train_golden_inputs(gd)
    fnet, mnet = vn.FeatureLearningNet(target), vn.MiddleConvNet()
    load_state(fnet, sd, ""); load_state(mnet, sd, "")
    fe = vn.VoxelNetFrontEnd(fnet.to(dev).train(), mnet.to(dev).train(), target)
    pts, offs, max_n = pack_frames([points])
    out = fe.forward_points_train(pts, offs, max_n)
    out.backward(dout.to(dev))
    stride = int(g["conv_stride"]); rows = []
    for mod in (fnet, mnet):
        for name, p in mod.named_parameters():
            if name.endswith("conv.bias"): continue
            want = g["grad:" + name].astype(np.float64).reshape(-1)
            got = p.grad.detach().double().cpu().numpy().reshape(-1)
            nrm = ""
            if name.endswith("conv.weight"):
                nrm = " norm_ratio %.4f" % (np.linalg.norm(got) / float(g["norm:" + name])); got = got[::stride]
            cos = float(got @ want / (np.linalg.norm(got) * np.linalg.norm(want)))
            devi = float(np.abs(got - want).max() / np.abs(want).max())
            rows.append("%-36s cos %.5f dev %.4f%s" % (name, cos, devi, nrm))
    print("run", rep); print("\n".join(rows))
