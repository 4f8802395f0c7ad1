"""CPU, container only: the oracle against the LIVE unmodified reference imported from
/root/reference (skipped where the reference is absent, e.g. on the GPU box)."""
import warnings

import numpy as np
import pytest
import torch

from oracle import model_torch as OM
from oracle import ref_import as R
from oracle import voxelize_np as OV

pytestmark = [
    pytest.mark.reference,
    pytest.mark.skipif(not R.reference_available(), reason="/root/reference not present"),
    pytest.mark.filterwarnings("ignore"),
]


@pytest.mark.parametrize("frame,target", [(0, "Car"), (57, "Car"), (107, "Pedestrian")])
def test_voxelizer_on_real_kitti_frames(frame, target):
    pcl = np.fromfile(R.kitti_frame_path(frame), dtype=np.float32).reshape(-1, 4)
    want = R.ref_pcl_to_voxels(pcl, target)
    got = OV.pcl_to_voxels_fast(pcl, target)
    for k in want:
        assert got[k].dtype == want[k].dtype and got[k].shape == want[k].shape, k
        assert np.array_equal(got[k].view(np.uint8), want[k].view(np.uint8)), k


def test_voxelizer_with_reference_shuffle_seeded():
    """With the real in-place shuffle on both sides and the same numpy seed."""
    ref = R.load_reference()
    pcl = np.fromfile(R.kitti_frame_path(3), dtype=np.float32).reshape(-1, 4)[::3].copy()
    a = pcl.copy()
    np.random.seed(7)
    want = ref.utils.pcl_to_voxels(a, "Car", False)
    b = pcl.copy()
    np.random.seed(7)
    np.random.shuffle(b)
    assert np.array_equal(a, b)
    got = OV.pcl_to_voxels_loop(b, "Car")
    for k in want:
        assert np.array_equal(got[k].view(np.uint8), want[k].view(np.uint8)), k


def test_prepare_voxel_matches_reference():
    ref = R.load_reference()
    pcl = np.fromfile(R.kitti_frame_path(1), dtype=np.float32).reshape(-1, 4)[::5]
    d = [OV.pcl_to_voxels_fast(pcl, "Car"), OV.pcl_to_voxels_fast(pcl[::2], "Car")]
    f0, n0, c0 = ref.dataset.prepare_voxel(d)
    f1, n1, c1 = OV.prepare_voxel(d)
    for a, b in zip(c0, c1):
        assert a.dtype == b.dtype and np.array_equal(a, b)


@pytest.mark.parametrize("target", ["Car", "Pedestrian"])
def test_model_oracle_matches_reference_modules(target):
    sd = OM.random_state_dict(31)
    pcl = np.fromfile(R.kitti_frame_path(2), dtype=np.float32).reshape(-1, 4)[::8]
    vox = OV.pcl_to_voxels_fast(pcl, target)
    feat = torch.from_numpy(vox["feature_buffer"][:200].copy())
    coord = torch.from_numpy(np.pad(vox["coordinate_buffer"][:200], ((0, 0), (1, 0)), constant_values=0))
    with R.ref_model_cfg(target) as ref, torch.no_grad():
        fnet = ref.model.FeatureLearningNet().eval()
        fnet.load_state_dict({k: v for k, v in sd.items() if k.startswith("vfe_")})
        want = fnet([feat], [coord])
        got = OM.feature_learning_net([feat], [coord], sd, target)
        assert got.shape == want.shape and torch.allclose(got, want, rtol=1e-6, atol=1e-6)
        mid = ref.model.MiddleConvNet().eval()
        mid.load_state_dict({k: v for k, v in sd.items() if k.startswith("middle_layer.")}, strict=False)
        x = torch.randn((1, 10, 16, 24, 128), generator=torch.Generator().manual_seed(1))
        want_mid = mid.middle_layer(x.permute(0, 4, 1, 2, 3)).reshape((1, -1, 16, 24))
        assert torch.allclose(OM.middle_layers(x, sd), want_mid, rtol=1e-5, atol=1e-5)


def test_training_mode_oracle_matches_reference_modules():
    """model.train(True) (train.py:148): batch-statistics BatchNorm in both nets, and the running statistics
    they leave behind."""
    sd = OM.random_state_dict(32)
    pcl = np.fromfile(R.kitti_frame_path(4), dtype=np.float32).reshape(-1, 4)[::8]
    vox = OV.pcl_to_voxels_fast(pcl, "Car")
    feat = torch.from_numpy(vox["feature_buffer"][:300].copy())
    coord = torch.from_numpy(np.pad(vox["coordinate_buffer"][:300], ((0, 0), (1, 0)), constant_values=0))
    with R.ref_model_cfg("Car") as ref, torch.no_grad():
        fnet = ref.model.FeatureLearningNet().train()
        fnet.load_state_dict({k: v for k, v in sd.items() if k.startswith("vfe_")})
        want = fnet([feat], [coord])
        new = {}
        got = OM.scatter_dense(OM.svfe_voxelwise(feat, sd, training=True, new_stats=new), coord, 1, "Car")
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-6)
        for k, v in new.items():
            assert torch.allclose(v, fnet.state_dict()[k], rtol=1e-6, atol=1e-7), k
        mid = ref.model.MiddleConvNet().train()
        mid.load_state_dict({k: v for k, v in sd.items() if k.startswith("middle_layer.")}, strict=False)
        x = torch.randn((2, 10, 16, 24, 128), generator=torch.Generator().manual_seed(2))
        want_mid = mid.middle_layer(x.permute(0, 4, 1, 2, 3)).reshape((2, -1, 16, 24))
        new = {}
        assert torch.allclose(OM.middle_layers(x, sd, training=True, new_stats=new), want_mid, rtol=1e-4, atol=1e-5)
        for k, v in new.items():
            assert torch.allclose(v, mid.state_dict()[k], rtol=1e-5, atol=1e-6), k


def test_state_dict_names_match_reference():
    ref = R.load_reference()
    import voxelnet_pytorch_b200 as vn

    rf = set(ref.model.FeatureLearningNet().state_dict().keys())
    assert rf == set(vn.FeatureLearningNet().state_dict().keys())
    rm = {k for k in ref.model.MiddleConvNet().state_dict().keys() if k.startswith("middle_layer.")}
    assert rm == set(vn.MiddleConvNet().state_dict().keys())
    assert rf | rm == set(OM.random_state_dict(0).keys())


def test_target_assignment_against_live_reference():
    """A second seed of label lines through the reference's generate_targets vs oracle + the product's host side."""
    from oracle import make_golden_targets as MG
    from oracle import targets_np as OT
    from voxelnet_pytorch_b200 import targets as T

    ref = R.load_reference()
    labels = MG.synthetic_labels(seed=5)[::2]  # the 10-line frame and the single-object frame
    lab = np.empty(len(labels), dtype=object)
    for i, f in enumerate(labels):
        lab[i] = f
    anchors = ref.utils.generate_anchors()
    assert np.array_equal(anchors, T.generate_anchors("Car"))
    want_boxes = ref.utils.label_to_gt_box_3d(lab, "Car", "lidar")
    boxes = T.label_to_gt_box_3d(labels, "Car", "lidar")
    assert all(np.array_equal(a, b) for a, b in zip(boxes, want_boxes))
    wp, wn, wt = ref.utils.generate_targets(lab, anchors.shape[:2], anchors)
    c = T.anchor_config("Car")
    gp, gn, gt = OT.generate_targets(boxes, anchors.shape[:2], anchors, c.rpn_pos_iou, c.rpn_neg_iou, c.anchor_h)
    assert np.array_equal(gp, wp) and np.array_equal(gn, wn) and np.array_equal(gt, wt) and wp.sum() > 0


def test_hybrid_rpn_wraps_the_live_reference_rpn():
    """HybridRPN takes the unmodified reference's MiddleConvNet: the attribute and parameter names it relies on exist, and
    it picks the layers the tcgen05 kernel covers (block1 / block2 stride-1 tails, deconv1).  Construction only - the
    kernels need a GPU (tests/test_gpu_rpn_conv2d.py)."""
    import voxelnet_pytorch_b200 as vn

    ref = R.load_reference()
    net = ref.model.MiddleConvNet().eval()
    rpn = vn.HybridRPN(net)
    assert rpn.block1.on_tensor_cores(352) == [True] * 5 and rpn.block2.on_tensor_cores(176) == [True] * 6
    assert rpn.block3.on_tensor_cores(88) == [True] * 6
    assert isinstance(rpn.deconv_fast[0], vn.DeConvMD2d) and isinstance(rpn.deconv_fast[1], vn.DeConvUpMD2d) \
        and isinstance(rpn.deconv_fast[2], vn.DeConvUpMD2d) and rpn.deconv_fast[2].k == 4
    assert set(rpn.block1.fast[1].state_dict()) == set(net.block1[1].state_dict())
    assert net.block1[0].conv.weight.dtype == torch.float32


def test_rpn_mirror_has_the_reference_parameter_names_and_output():
    """vn.RPN is the RPN half of the reference's MiddleConvNet under the same parameter names: the reference's state dict
    (minus middle_layer.*) loads strictly, and the two forwards agree bit for bit on a small map (stock fp32 layers on both
    sides)."""
    import voxelnet_pytorch_b200 as vn

    ref = R.load_reference()
    torch.manual_seed(3)
    net = ref.model.MiddleConvNet().eval()
    for m in net.modules():
        if isinstance(m, torch.nn.BatchNorm2d):
            m.running_mean.normal_(0, 0.1)
            m.running_var.uniform_(0.5, 1.5)
    mine = vn.RPN("Car").eval()
    sd = {k: v for k, v in net.state_dict().items() if not k.startswith("middle_layer.")}
    assert set(sd) == set(mine.state_dict())
    mine.load_state_dict(sd, strict=True)
    assert mine.output_shape == list(net.output_shape)

    x = torch.randn(1, 128, 32, 16)
    with torch.no_grad():
        h = net.block1(x)
        d1 = net.deconv1(h)
        h = net.block2(h)
        d2 = net.deconv2(h)
        h = net.block3(h)
        d3 = net.deconv3(h)
        cat = torch.cat([d3, d2, d1], 1)
        want = torch.sigmoid(net.prob_conv(cat)), net.reg_conv(cat)
        got = mine(x)
    assert torch.equal(got[0], want[0]) and torch.equal(got[1], want[1])
    assert vn.RPN("Pedestrian").block1[0].conv.stride == (1, 1)


def test_middle_rpn_is_a_strict_state_dict_match_of_the_reference_middle_conv_net():
    """vn.MiddleRPN = the reference's whole MiddleConvNet under its names: same state-dict keys in the same order, strict
    load, same output_shape, and the RPN half of its forward (stock fp32) agrees bit for bit on CPU."""
    import voxelnet_pytorch_b200 as vn

    ref = R.load_reference()
    torch.manual_seed(4)
    net = ref.model.MiddleConvNet().eval()
    mine = vn.MiddleRPN("Car").eval()
    assert list(mine.state_dict()) == list(net.state_dict())
    mine.load_state_dict(net.state_dict(), strict=True)
    assert mine.output_shape == list(net.output_shape)
    x = torch.randn(1, 128, 16, 16)
    with torch.no_grad():
        got = vn.RPN.forward(mine, x)
        h = net.block1(x)
        d1 = net.deconv1(h)
        h = net.block2(h)
        d2 = net.deconv2(h)
        h = net.block3(h)
        cat = torch.cat([net.deconv3(h), d2, d1], 1)
        want = torch.sigmoid(net.prob_conv(cat)), net.reg_conv(cat)
    assert torch.equal(got[0], want[0]) and torch.equal(got[1], want[1])
    assert isinstance(mine.hybrid(), vn.HybridRPN)


def test_rpn_golden_is_what_the_live_reference_computes():
    """tests/golden/rpn_car.npz regenerates bit for bit from the live reference (same torch build as the generator's)."""
    import hashlib

    from tests import golden_util as G

    g = G.load("rpn_car.npz")
    ref = R.load_reference()
    net = ref.model.MiddleConvNet().eval()
    assert OM.fill_rpn_parameters(net, int(g["seed"])) == str(g["params_sha256"])
    x = OM.rpn_golden_input(int(g["seed"]), tuple(int(v) for v in g["input_shape"]))
    assert hashlib.sha256(x.numpy().tobytes()).hexdigest() == str(g["input_sha256"])
    with torch.no_grad():
        h = net.block1(x)
        d1 = net.deconv1(h)
        h = net.block2(h)
        d2 = net.deconv2(h)
        h = net.block3(h)
        cat = torch.cat([net.deconv3(h), d2, d1], 1)
        prob, reg = torch.sigmoid(net.prob_conv(cat)), net.reg_conv(cat)
    assert np.array_equal(prob.numpy(), g["prob"]) and np.array_equal(reg.numpy(), g["reg"])
