"""GPU parity of the training-path backward kernels against torch-CPU autograd over the oracle
(the reference trains with loss.backward(), train.py:149, through model.py:74-82, 95-106, 158-167, 257-262)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import model_torch as OM
from tests.helpers import load_state, oracle_batch, rel_max_err, synth_frames

pytestmark = pytest.mark.gpu


def _bf16_exact(shape, gen, scale=1.0):
    return (torch.randn(shape, generator=gen) * scale).to(torch.bfloat16).float()


@pytest.mark.parametrize("final", [False, True])
def test_bn_relu_bwd_matches_autograd(cuda_device, final):
    from voxelnet_pytorch_b200 import ops

    g = torch.Generator().manual_seed(60)
    B, D, H, W = 2, 2, 9, 40
    raw = _bf16_exact((B, D, H, W, 64), g).requires_grad_(True)
    gamma = (0.5 + torch.rand(64, generator=g)).requires_grad_(True)
    beta = ((torch.rand(64, generator=g) - 0.5) * 0.6).requires_grad_(True)
    x = raw.permute(0, 4, 1, 2, 3)
    y = F.relu(F.batch_norm(x, None, None, gamma, beta, training=True, eps=OM.BN_EPS))
    if final:
        y = y.reshape(B, -1, H, W)
        dy = torch.randn(y.shape, generator=g)
    else:
        y = y.permute(0, 2, 3, 4, 1)
        dy = _bf16_exact(y.shape, g)
    y.backward(dy)
    with torch.no_grad():
        mean = raw.mean(dim=(0, 1, 2, 3))
        var = raw.var(dim=(0, 1, 2, 3), unbiased=False)
        invstd = torch.rsqrt(var + OM.BN_EPS)
        scale = gamma * invstd
        shift = beta - mean * scale
    dev = cuda_device
    dy_dev = dy.to(dev) if final else dy.to(dev).to(torch.bfloat16)
    draw, dgamma, dbeta = ops.bn_relu_bwd(raw.detach().to(dev).to(torch.bfloat16), dy_dev.contiguous(), scale.to(dev), shift.to(dev),
                                          mean.to(dev), invstd.to(dev), final=final)
    assert rel_max_err(draw.float().cpu(), raw.grad) < 6e-3  # bf16 storage of draw
    assert rel_max_err(dgamma.cpu(), gamma.grad) < 1e-4
    assert rel_max_err(dbeta.cpu(), beta.grad) < 1e-4


def test_conv1_dgrad_gather_matches_autograd(cuda_device):
    """Data gradient of the first middle layer at the occupied cells (to_dense backward fused in)."""
    from voxelnet_pytorch_b200 import ops

    g = torch.Generator().manual_seed(61)
    B, D, H, W = 2, 10, 12, 16
    Dr = 5
    w = _bf16_exact((64, 128, 3, 3, 3), g, 0.05)
    draw = _bf16_exact((B, Dr, H, W, 64), g)
    x = torch.zeros((B, 128, D, H, W), requires_grad=True)
    y = F.conv3d(x, w, None, stride=(2, 1, 1), padding=(1, 1, 1))
    y.backward(draw.permute(0, 4, 1, 2, 3))
    occ = torch.rand((B, D, H, W), generator=g) < 0.08
    occ[0, 0, 0, 0] = occ[1, 9, H - 1, W - 1] = occ[0, 9, 0, W - 1] = True  # corners and both depth borders
    coords = occ.nonzero().contiguous()  # lexicographic (b, z, y, x), as np.unique orders them
    want = x.grad.permute(0, 2, 3, 4, 1)[occ]
    got = ops.conv1_dgrad_gather(draw.to(cuda_device).to(torch.bfloat16), ops.conv1_weight_taps(w.to(cuda_device)),
                                 coords.to(cuda_device), 2, 1).cpu()
    assert got.shape == want.shape
    assert rel_max_err(got, want) < 1e-5


def _rel_l2(got, want):
    return float((got.double() - want.double()).norm() / want.double().norm().clamp_min(1e-30))


def _svfe_oracle_grads(feature, sd, dvw):
    keys = [k for k in sd if k.startswith("vfe_") and (k.endswith("weight") or k.endswith("bias"))]
    leaf = {k: (sd[k].clone().requires_grad_(True) if k in keys else sd[k]) for k in sd}
    vw = OM.svfe_voxelwise(feature, leaf, training=True)
    vw.backward(dvw)
    return vw.detach(), {k: leaf[k].grad for k in keys}


def test_svfe_backward_matches_autograd(cuda_device):
    """Gradients of both VFE layers (Linear weight/bias, BatchNorm1d weight/bias) with batch statistics over
    all K*T rows; voxels with one point, with > T points, and pad rows that win the unmasked max."""
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(31)
    frames = [f[::3].copy() for f in synth_frames(2, 140)]
    frames[1] = np.concatenate([frames[1], np.tile(np.array([[20.05, 1.05, -1.0, 0.5]], np.float32), (80, 1))])
    _, feats, _, _ = oracle_batch(frames, "Car")
    feature = torch.from_numpy(np.concatenate(feats))
    g = torch.Generator().manual_seed(32)
    dvw = torch.randn((feature.shape[0], 128), generator=g)
    want_vw, want = _svfe_oracle_grads(feature, sd, dvw)

    fnet = vn.FeatureLearningNet("Car")
    load_state(fnet, sd, "")
    fnet = fnet.to(cuda_device).train()
    vw = fnet.voxelwise(feature.to(cuda_device))
    assert vw.requires_grad
    assert torch.allclose(vw.detach().cpu(), want_vw, rtol=1e-4, atol=1e-4)  # |vw| reaches ~10
    vw.backward(dvw.to(cuda_device))
    got = {k: p.grad.cpu() for k, p in fnet.named_parameters()}
    assert set(got) == set(want)
    for k in sorted(want):
        err = rel_max_err(got[k], want[k])
        assert err < 2e-3, (k, err)


def test_middle_net_backward_matches_autograd(cuda_device):
    """Three ConvMD layers in training mode on a small grid: parameter gradients and the gradient of the
    voxel-wise features (through the sparse side channel) against fp32 autograd."""
    import voxelnet_pytorch_b200 as vn
    from voxelnet_pytorch_b200 import ops

    sd = OM.random_state_dict(33)
    g = torch.Generator().manual_seed(34)
    B, D, H, W = 2, 10, 40, 32
    occ = torch.rand((B, D, H, W), generator=g) < 0.03
    coords = occ.nonzero().contiguous()
    K = coords.shape[0]
    vw0 = torch.randn((K, 128), generator=g)
    dout = torch.randn((B, 128, H, W), generator=g)

    keys = [k for k in sd if k.startswith("middle_layer.") and (k.endswith("weight") or k.endswith("bias"))]
    leaf = {k: (sd[k].clone().requires_grad_(True) if k in keys else sd[k]) for k in sd}
    vw_ref = vw0.clone().requires_grad_(True)
    dense_ref = torch.zeros((B, D, H, W, 128)).index_put((coords[:, 0], coords[:, 1], coords[:, 2], coords[:, 3]), vw_ref)
    # bf16-emulating oracle: same rounding points as the CUDA path, so the ReLU masks agree (see its docstring)
    out_ref = OM.middle_layers_bf16_emulated(dense_ref, leaf)
    out_ref.backward(dout)

    mnet = vn.MiddleConvNet()
    load_state(mnet, sd, "")
    mnet = mnet.to(cuda_device).train()
    vw = vw0.to(cuda_device).requires_grad_(True)
    dense = torch.zeros((B, D, H, W, 128), dtype=torch.bfloat16, device=cuda_device)
    ops.scatter_dense_bf16(vw.detach(), coords.to(cuda_device), dense)
    dense._vn_sparse = (vw, coords.to(cuda_device))
    out = mnet(dense)
    assert rel_max_err(out.detach().cpu(), out_ref.detach()) < 1e-2
    assert rel_max_err(out.detach().cpu(), OM.middle_layers(dense_ref.detach(), sd, training=True)) < 1e-2  # plain fp32
    out.backward(dout.to(cuda_device))
    errs = {"vw": rel_max_err(vw.grad.cpu(), vw_ref.grad)}
    l2 = {"vw": _rel_l2(vw.grad.cpu(), vw_ref.grad)}
    for name, p in mnet.named_parameters():
        k = "middle_layer." + name[len("middle_layer."):] if name.startswith("middle_layer.") else name
        want = leaf[k].grad
        if k.endswith("conv.bias"):  # exactly zero under batch-statistics BN; autograd leaves rounding noise
            assert float(p.grad.abs().max()) == 0.0
            assert float(want.abs().max()) < 1e-3 * float(leaf[k.replace("bias", "weight")].grad.abs().max())
            continue
        errs[k] = rel_max_err(p.grad.cpu(), want)
        l2[k] = _rel_l2(p.grad.cpu(), want)
    # what is left are single ReLU flips (accumulation order) and the bf16 storage of the gradients between layers
    bad = {k: v for k, v in errs.items() if not v < 8e-2}
    assert not bad, errs
    bad = {k: v for k, v in l2.items() if not v < 3e-2}
    assert not bad, l2


def test_front_end_training_step_end_to_end(cuda_device):
    """points -> voxelize -> SVFE -> grid -> 3 x ConvMD in training mode through VoxelNetFrontEnd, backward
    from a random upstream gradient: every hot-path parameter gradient against the fp32 CPU oracle."""
    import voxelnet_pytorch_b200 as vn
    from voxelnet_pytorch_b200.utils import pack_frames

    target = "Pedestrian"
    sd = OM.random_state_dict(35)
    frames = [f[::2].copy() for f in synth_frames(1, 150)]
    _, feats, _, coords = oracle_batch(frames, target)
    keys = [k for k in sd if k.endswith("weight") or k.endswith("bias")]
    leaf = {k: (sd[k].clone().requires_grad_(True) if k in keys else sd[k]) for k in sd}
    vw_ref = OM.svfe_voxelwise(torch.from_numpy(np.concatenate(feats)), leaf, training=True)
    c = torch.from_numpy(np.concatenate(coords))
    D, H, W, _ = OM.grid_dims(target)
    dense_ref = torch.zeros((1, D, H, W, 128)).index_put((c[:, 0], c[:, 1], c[:, 2], c[:, 3]), vw_ref)
    out_ref = OM.middle_layers_bf16_emulated(dense_ref, leaf)
    g = torch.Generator().manual_seed(36)
    dout = torch.randn(out_ref.shape, generator=g)
    out_ref.backward(dout)

    fnet, mnet = vn.FeatureLearningNet(target), vn.MiddleConvNet()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    fe = vn.VoxelNetFrontEnd(fnet.to(cuda_device).train(), mnet.to(cuda_device).train(), target)
    points, offsets, max_n = pack_frames(frames)
    out = fe.forward_points_train(points, offsets, max_n)
    assert rel_max_err(out.detach().cpu(), out_ref.detach()) < 1e-2
    out.backward(dout.to(cuda_device))
    errs = {}
    for mod in (fnet, mnet):
        for name, p in mod.named_parameters():
            if name.endswith("conv.bias"):
                continue
            errs[name] = (rel_max_err(p.grad.cpu(), leaf[name].grad), _rel_l2(p.grad.cpu(), leaf[name].grad))
    bad = {k: v for k, v in errs.items() if not (v[0] < 8e-2 and v[1] < 6e-2)}
    assert not bad, errs
    # a second step reuses the resident grid: the cells of step 1 are cleared first
    for mod in (fnet, mnet):
        mod.zero_grad()
    out2 = fe.forward_points_train(points, offsets, max_n)
    assert rel_max_err(out2.detach().cpu(), out_ref.detach()) < 1e-2
    # back to eval on the same front end: BN is re-folded from the running statistics the two steps updated
    fnet.eval(), mnet.eval()
    got_eval = fe.forward_points(points, offsets, max_n).clone()
    fresh = vn.VoxelNetFrontEnd(fnet, mnet, target)
    assert torch.equal(got_eval, fresh.forward_points(points, offsets, max_n))
    assert int(fnet.vfe_1.bn.num_batches_tracked) == 2 and int(mnet.middle_layer[2].batch_norm.num_batches_tracked) == 2


def test_training_step_with_an_empty_frame_and_through_the_host_stream(cuda_device):
    """Ragged batch: one frame without a single point in range next to a normal one; fed through the pinned
    host stream.  The step equals the one-frame-at-index-1 oracle forward and leaves finite gradients; a batch
    with no voxel at all raises (as BatchNorm does on an empty batch)."""
    import voxelnet_pytorch_b200 as vn

    target = "Pedestrian"
    sd = OM.random_state_dict(37)
    full = synth_frames(1, 160)[0][::4].copy()
    empty = np.array([[500.0, 500.0, 50.0, 0.1]], np.float32)  # out of range: cropped
    _, feats, _, coords = oracle_batch([empty, full], target)
    assert feats[0].shape[0] == 0
    with torch.no_grad():
        vw = OM.svfe_voxelwise(torch.from_numpy(feats[1]), sd, training=True)
        ref = OM.middle_layers(OM.scatter_dense(vw, torch.from_numpy(coords[1]), 2, target), sd, training=True)
    fnet, mnet = vn.FeatureLearningNet(target), vn.MiddleConvNet()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    fe = vn.VoxelNetFrontEnd(fnet.to(cuda_device).train(), mnet.to(cuda_device).train(), target)
    hb = vn.PinnedFrameBatch.from_frames([empty, full])
    stream = vn.HostFrameStream(fe, 2, hb.n_points)
    stream.upload(hb)
    out = stream.forward()
    assert stream.voxel_counts_host().tolist() == [0, 0, feats[1].shape[0]]
    assert rel_max_err(out.detach().cpu(), ref) < 1e-2
    out.backward(torch.ones_like(out) * 1e-3)
    for mod in (fnet, mnet):
        for name, p in mod.named_parameters():
            assert bool(torch.isfinite(p.grad).all()), name
    hb0 = vn.PinnedFrameBatch.from_frames([empty, empty])
    stream.upload(hb0)
    with pytest.raises(RuntimeError):
        stream.forward()
