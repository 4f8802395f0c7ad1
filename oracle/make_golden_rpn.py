"""Generate tests/golden/rpn_car.npz by running the RPN half of the UNMODIFIED reference's MiddleConvNet (container only; the
generating script of the committed fixture).

    python -m oracle.make_golden_rpn

What runs: `MiddleConvNet()` of voxelnet/model.py:202-254 (Car config: block1 starts with a stride-2 layer), eval mode, its
RPN parameters filled by `oracle.model_torch.fill_rpn_parameters` (name-keyed, seeded: the mirror on the GPU box fills its
own copy the same way instead of shipping 18 MB of weights), and the tail of its forward (model.py:263-281: block1,
deconv1, block2, deconv2, block3, deconv3, concat, the two 1x1 heads, sigmoid) on `rpn_golden_input`.  Stored: both
outputs in full (49 KB), the sha256 of the filled parameters and of the input, and the seed.
"""
from __future__ import annotations

import hashlib
import os
import sys
import warnings

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import model_torch as OM  # noqa: E402
from oracle import ref_import as R  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")
SEED = 5


def main() -> None:
    warnings.filterwarnings("ignore")
    ref = R.load_reference()
    torch.manual_seed(0)
    net = ref.model.MiddleConvNet().eval()
    params_sha = OM.fill_rpn_parameters(net, SEED)
    x = OM.rpn_golden_input(SEED)
    with torch.no_grad():
        h = net.block1(x)
        d1 = net.deconv1(h)
        h = net.block2(h)
        d2 = net.deconv2(h)
        h = net.block3(h)
        d3 = net.deconv3(h)
        cat = torch.cat([d3, d2, d1], dim=1)
        prob, reg = torch.sigmoid(net.prob_conv(cat)), net.reg_conv(cat)
    out = os.path.join(GOLDEN, "rpn_car.npz")
    np.savez_compressed(
        out, seed=np.int64(SEED), prob=prob.numpy(), reg=reg.numpy(), params_sha256=np.str_(params_sha),
        input_sha256=np.str_(hashlib.sha256(x.numpy().tobytes()).hexdigest()), input_shape=np.array(x.shape, dtype=np.int64),
        cat_abs_mean=np.float64(cat.abs().mean()), torch_version=np.str_(torch.__version__))
    print("wrote", out, os.path.getsize(out), "bytes; prob", tuple(prob.shape), float(prob.min()), float(prob.max()), "reg |max|",
          float(reg.abs().max()), "cat |mean|", float(cat.abs().mean()))


if __name__ == "__main__":
    main()
