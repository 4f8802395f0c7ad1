"""CPU: the oracle against the golden vectors frozen from the unmodified reference."""
import os

import numpy as np
import pytest
import torch

from oracle import model_torch as OM
from oracle import voxelize_np as OV
from tests import golden_util as G


@pytest.mark.parametrize("case", range(5))
def test_voxelizer_oracle_fast_matches_reference_golden(case):
    name, target, pts = G.voxel_cases()[case]
    G.check_voxel_dict_against_golden(OV.pcl_to_voxels_fast(pts, target), name)


def test_voxelizer_oracle_loop_matches_reference_golden():
    name, target, pts = G.voxel_cases()[0]
    G.check_voxel_dict_against_golden(OV.pcl_to_voxels_loop(pts, target), name)


def test_loop_and_fast_agree_on_edge_cases():
    rng = np.random.default_rng(0)
    pts = np.concatenate([
        np.tile(np.array([[10.05, 0.05, -1.0, 0.5]], np.float32), (80, 1)) + rng.uniform(0, 0.09, (80, 4)).astype(np.float32),
        np.array([[0.0, -40.0, -3.0, 0.1], [70.4, 0, 0, 0.2], [-0.0, 0.0, -0.0, 0.0], [0, 0, 0, 0], [-1e-9, 0, 0, 0.5]], np.float32),
    ])
    for target in ("Car", "Cyclist"):
        a, b = OV.pcl_to_voxels_loop(pts, target), OV.pcl_to_voxels_fast(pts, target)
        for k in a:
            assert a[k].dtype == b[k].dtype and np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8)), (target, k)
    e = OV.pcl_to_voxels_fast(np.zeros((0, 4), np.float32), "Car")
    assert e["feature_buffer"].shape == (0, 35, 7) and e["coordinate_buffer"].shape == (0, 3)


def test_prepare_voxel_batch_column():
    pts = G.kitti_sub_points()
    d = [OV.pcl_to_voxels_fast(pts[:5000], "Car"), OV.pcl_to_voxels_fast(pts[5000:9000], "Car")]
    f, n, c = OV.prepare_voxel(d)
    assert c[1].shape[1] == 4 and c[1].dtype == np.int64 and (c[1][:, 0] == 1).all() and (c[0][:, 0] == 0).all()
    assert np.array_equal(c[1][:, 1:], d[1]["coordinate_buffer"])


def _model_inputs():
    g = G.load("model_ref_outputs.npz")
    sd = OM.random_state_dict(int(g["state_seed"]))
    K = int(g["K"])
    vox = OV.pcl_to_voxels_fast(G.kitti_sub_points(), "Car")
    feat = torch.from_numpy(vox["feature_buffer"][:K].copy())
    coord = torch.from_numpy(np.pad(vox["coordinate_buffer"][:K], ((0, 0), (1, 0)), constant_values=0))
    return g, sd, feat, coord


def test_svfe_oracle_matches_reference_golden():
    g, sd, feat, coord = _model_inputs()
    with torch.no_grad():
        vw = OM.svfe_voxelwise(feat, sd)
        assert torch.allclose(vw, torch.from_numpy(g["voxelwise"]), rtol=1e-6, atol=1e-6)
        mask = OM.feature_mask(feat)
        v1 = OM.vfe_layer(feat, mask, sd, "vfe_1.")
        assert torch.allclose(v1, torch.from_numpy(g["vfe1_out"]), rtol=1e-6, atol=1e-6)
        gen = torch.Generator().manual_seed(int(g["vfe2_seed"]))
        x2 = torch.randn((64, 35, 32), generator=gen)
        m2 = torch.rand((64, 35, 1), generator=gen) > 0.3
        v2 = OM.vfe_layer(x2, m2, sd, "vfe_2.")
        assert torch.allclose(v2, torch.from_numpy(g["vfe2_out"]), rtol=1e-6, atol=1e-6)
        dense = OM.feature_learning_net([feat], [coord], sd, "Car")
        assert dense.shape == (1, 10, 400, 352, 128)
        assert int((dense.abs().sum(-1) != 0).sum()) == int(g["dense_nonzero_rows"])


def test_middle_oracle_matches_reference_golden():
    g = G.load("model_ref_outputs.npz")
    sd = OM.random_state_dict(int(g["state_seed"]))
    B, D, H, W = (int(v) for v in g["middle_shape"])
    gen = torch.Generator().manual_seed(int(g["middle_seed"]))
    xd = torch.randn((B, D, H, W, 128), generator=gen) * (torch.rand((B, D, H, W, 1), generator=gen) > 0.8)
    with torch.no_grad():
        y = OM.middle_layers(xd, sd)
        y1 = OM.middle_layers(xd, sd, upto=1)
    assert torch.allclose(y1, torch.from_numpy(g["middle_layer0_out"]), rtol=1e-5, atol=1e-5)
    assert torch.allclose(y, torch.from_numpy(g["middle_out"]), rtol=1e-5, atol=1e-5)


def _train_golden_inputs(golden_dir):
    """Inputs of tests/golden/train_grads_ped.npz (made by oracle/make_golden_train.py from the live reference)."""
    import torch

    from oracle.make_golden_train import upstream_gradient

    g = np.load(os.path.join(golden_dir, "train_grads_ped.npz"))
    points = np.load(os.path.join(golden_dir, "voxelize_kitti0_sub_car.npz"))["points"]
    target = str(g["target"])
    D, H, W, _ = OM.grid_dims(target)
    dout = upstream_gradient((1, 128, H, W), int(g["dout_seed"]))
    return g, points, target, OM.random_state_dict(int(g["state_seed"])), dout


def test_oracle_training_gradients_match_reference_golden(golden_dir):
    """Autograd through the oracle in training mode reproduces the gradients the UNMODIFIED reference produced
    (train.py:148-149 semantics: batch-statistics BN in both VFE layers and the three ConvMD layers)."""
    import torch

    g, points, target, sd, dout = _train_golden_inputs(golden_dir)
    v = OV.pcl_to_voxels_fast(points, target)
    assert len(v["number_buffer"]) == int(g["n_voxels"])
    keys = [k for k in sd if k.endswith("weight") or k.endswith("bias")]
    leaf = {k: (sd[k].clone().requires_grad_(True) if k in keys else sd[k]) for k in sd}
    feats, _, coords = OV.prepare_voxel([v])
    vw = OM.svfe_voxelwise(torch.from_numpy(feats[0]), leaf, training=True)
    dense = OM.scatter_dense(vw, torch.from_numpy(coords[0]), 1, target)
    out = OM.middle_layers(dense, leaf, training=True)
    assert abs(float(out.detach().abs().max()) - float(g["out_abs_max"])) < 1e-4 * float(g["out_abs_max"])
    # (fp32 on both sides, but the golden was made on another host: the conv primitive's summation order differs between
    # CPUs, and three batch-statistics layers amplify it - 1e-3 of max here, 1e-4 in the container that made the golden)
    assert np.allclose(out.detach()[0, ::16, ::8, ::8].numpy(), g["out_sample"], rtol=1e-3, atol=1e-3 * float(g["out_abs_max"]))
    out.backward(dout)
    stride = int(g["conv_stride"])
    for k in keys:
        want = g["grad:" + k]
        got = leaf[k].grad.numpy()
        if k.endswith("conv.weight"):
            assert abs(float(np.linalg.norm(got)) - float(g["norm:" + k])) < 1e-3 * float(g["norm:" + k]), k
            got = got.reshape(-1)[::stride]
        scale = max(float(np.abs(want).max()), 1e-20)
        if k.endswith("conv.bias"):  # rounding noise around zero in both
            assert float(np.abs(got).max()) < 1e-2 * float(np.abs(g["grad:" + k.replace("bias", "weight")]).max())
            continue
        assert float(np.abs(got - want).max()) / scale < 5e-3, (k, float(np.abs(got - want).max()) / scale)


def test_oracle_whole_chain_matches_reference_golden_on_a_full_kitti_frame():
    """CPU: the oracle chain on the committed full KITTI frame against the golden the UNMODIFIED reference produced
    (oracle/make_golden_chain.py).  fp32 on both sides; only the conv primitive's summation order may differ
    between hosts, hence 1e-5 of max instead of bit equality."""
    import os

    import torch

    from oracle import cpu_chain

    g = G.load("chain_kitti0_car.npz")
    pcl = np.fromfile(os.path.join(G.GOLDEN, "kitti_0000000000.bin"), dtype=np.float32).reshape(-1, 4)
    v = OV.pcl_to_voxels_fast(pcl, "Car")
    assert G.sha(v["coordinate_buffer"]) == str(g["coordinate_sha256"])
    assert G.sha(v["number_buffer"]) == str(g["number_sha256"])
    assert G.sha(v["feature_buffer"]) == str(g["feature_sha256"])
    out, _ = cpu_chain.front_end_cpu([pcl], OM.random_state_dict(int(g["state_seed"])), "Car", faithful_voxelizer=False)
    assert tuple(out.shape) == tuple(g["out_shape"])
    err = float((out[:, :, 1::4, 2::4] - torch.from_numpy(g["out_sample"])).abs().max()) / float(g["out_max_abs"])
    assert err < 1e-5, err
    sums = out.double().sum(dim=(0, 2, 3)).numpy()
    assert np.allclose(sums, g["out_channel_sums"], rtol=1e-6, atol=1e-3)


def test_rpn_mirror_matches_the_reference_rpn_golden():
    """tests/golden/rpn_car.npz = the RPN half of the unmodified reference's MiddleConvNet (oracle/make_golden_rpn.py) on a
    seeded map with name-keyed seeded parameters.  The stock fp32 mirror `vn.RPN('Car')`, filled the same way, reproduces it
    on the CPU (float32 convolutions on whatever host kernels: 1e-4 of the output range)."""
    import hashlib

    import voxelnet_pytorch_b200 as vn

    g = G.load("rpn_car.npz")
    seed = int(g["seed"])
    net = vn.RPN("Car").eval()
    assert OM.fill_rpn_parameters(net, seed) == str(g["params_sha256"])
    x = OM.rpn_golden_input(seed, tuple(int(v) for v in g["input_shape"]))
    assert hashlib.sha256(x.numpy().tobytes()).hexdigest() == str(g["input_sha256"])
    with torch.no_grad():
        prob, reg = net(x)
    want_p, want_r = torch.from_numpy(g["prob"]), torch.from_numpy(g["reg"])
    assert prob.shape == want_p.shape and reg.shape == want_r.shape
    assert float((prob - want_p).abs().max()) <= 1e-4
    assert float((reg - want_r).abs().max()) <= 1e-4 * float(want_r.abs().max())
    # the whole-MiddleConvNet drop-in holds the same RPN under the same names
    both = vn.MiddleRPN("Car").eval()
    assert OM.fill_rpn_parameters(both, seed) == str(g["params_sha256"])
    with torch.no_grad():
        p2, r2 = vn.RPN.forward(both, x)
    assert torch.equal(p2, prob) and torch.equal(r2, reg)
