"""Debug aid: per-tensor gradient errors of the training backward vs the CPU oracle (run on the GPU box)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from oracle import model_torch as OM
from tests.helpers import load_state, oracle_batch, rel_max_err, synth_frames
from tests import test_gpu_train_backward as T
import voxelnet_pytorch_b200 as vn
from voxelnet_pytorch_b200 import ops

dev = torch.device("cuda", 0)

def svfe():
    sd = OM.random_state_dict(31)
    frames = [f[::3].copy() for f in synth_frames(2, 140)]
    frames[1] = np.concatenate([frames[1], np.tile(np.array([[20.05, 1.05, -1.0, 0.5]], np.float32), (80, 1))])
    _, feats, _, _ = oracle_batch(frames, "Car")
    feature = torch.from_numpy(np.concatenate(feats))
    g = torch.Generator().manual_seed(32)
    for variant in ("lo", "hi", "both"):
        dvw = torch.randn((feature.shape[0], 128), generator=g)
        if variant == "lo": dvw[:, 64:] = 0
        if variant == "hi": dvw[:, :64] = 0
        want_vw, want = T._svfe_oracle_grads(feature, sd, dvw)
        fnet = vn.FeatureLearningNet("Car"); load_state(fnet, sd, ""); fnet = fnet.to(dev).train()
        vw = fnet.voxelwise(feature.to(dev)); vw.backward(dvw.to(dev))
        print("svfe", variant, {k: "%.2e" % rel_max_err(p.grad.cpu(), want[k]) for k, p in fnet.named_parameters()})

def middle():
    sd = OM.random_state_dict(33)
    g = torch.Generator().manual_seed(34)
    B, D, H, W = 2, 10, 40, 32
    occ = torch.rand((B, D, H, W), generator=g) < 0.03
    coords = occ.nonzero().contiguous(); K = coords.shape[0]
    vw0 = torch.randn((K, 128), generator=g); dout = torch.randn((B, 128, H, W), generator=g)
    keys = [k for k in sd if k.startswith("middle_layer.") and (k.endswith("weight") or k.endswith("bias"))]
    leaf = {k: (sd[k].clone().requires_grad_(True) if k in keys else sd[k]) for k in sd}
    vw_ref = vw0.clone().requires_grad_(True)
    dense_ref = torch.zeros((B, D, H, W, 128)).index_put((coords[:, 0], coords[:, 1], coords[:, 2], coords[:, 3]), vw_ref)
    # keep intermediates
    out_ref, acts = OM.middle_layers_bf16_emulated(dense_ref, leaf, return_acts=True); out_ref.backward(dout)
    mnet = vn.MiddleConvNet(); load_state(mnet, sd, ""); mnet = mnet.to(dev).train()
    vw = vw0.to(dev).requires_grad_(True)
    dense = torch.zeros((B, D, H, W, 128), dtype=torch.bfloat16, device=dev)
    ops.scatter_dense_bf16(vw.detach(), coords.to(dev), dense); dense._vn_sparse = (vw, coords.to(dev))
    x = mnet.middle_layer[0].forward_ndhwc(dense, sparse=dense._vn_sparse); x.retain_grad(); a0 = x
    x = mnet.middle_layer[1].forward_ndhwc(x); x.retain_grad(); a1 = x
    out = mnet.middle_layer[2].forward_ndhwc(x, final=True)
    print("fwd", rel_max_err(out.detach().cpu(), out_ref.detach()))
    out.backward(dout.to(dev))
    print("dact1", rel_max_err(a1.grad.float().cpu().permute(0, 4, 1, 2, 3), acts[1].grad))
    print("dact0", rel_max_err(a0.grad.float().cpu().permute(0, 4, 1, 2, 3), acts[0].grad))
    print("vw", rel_max_err(vw.grad.cpu(), vw_ref.grad))
    for name, p in mnet.named_parameters():
        print(name, "%.2e" % rel_max_err(p.grad.cpu(), leaf[name].grad), float(p.grad.abs().max()), float(leaf[name].grad.abs().max()))

if __name__ == "__main__":
    svfe(); middle()
