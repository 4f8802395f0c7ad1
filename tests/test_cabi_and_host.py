"""CPU: the C-ABI library loads and exports every symbol include/voxelnet_b200.h declares; host-side
logic (config, batch packer, argument validation, no-GPU failure mode)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "voxelnet_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vn_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from voxelnet_pytorch_b200 import _cabi

    lib = _cabi.lib()
    names = _header_symbols()
    assert len(names) >= 19
    for n in names:
        assert hasattr(lib, n), "libvoxelnet_b200.so lacks %s" % n
    assert sorted(_cabi.PROTOTYPES) == names, "ctypes prototypes out of sync with the header"
    assert lib.vn_abi_version() == _cabi.VN_ABI_VERSION


def test_grid_constants_match_reference_values():
    from voxelnet_pytorch_b200 import _cabi, grid_config

    car, ped = _cabi.grid_for_target("Car"), _cabi.grid_for_target("Cyclist")
    assert (car.D, car.H, car.W, car.T) == (10, 400, 352, 35)
    assert (ped.D, ped.H, ped.W, ped.T) == (10, 200, 240, 45)
    assert (car.off_x, car.off_y, car.off_z) == (0.0, 40.0, 3.0) and ped.off_y == 20.0
    assert np.float32(car.vs_x) == np.float32(0.2) and np.float32(car.vs_z) == np.float32(0.4)
    for t, g in (("Car", car), ("Pedestrian", ped)):
        c = grid_config(t)
        assert (c.depth, c.height, c.width, c.points_per_voxel) == (g.D, g.H, g.W, g.T)
        # config.py:46-48: cells = range / voxel size
        assert c.width == int(round((c.x_range[1] - c.x_range[0]) / 0.2)) and c.height == int(round((c.y_range[1] - c.y_range[0]) / 0.2))


def test_bad_arguments_are_rejected_without_a_gpu():
    from voxelnet_pytorch_b200 import _cabi

    lib = _cabi.lib()
    g = _cabi.grid_for_target("Car")
    assert lib.vn_voxelize_workspace_bytes(ctypes.byref(g), 16, 2_000_000) > 0
    assert lib.vn_voxelize_workspace_bytes(ctypes.byref(g), 0, 10) == 0
    rc = lib.vn_voxelize_index(ctypes.byref(g), None, None, 1, 10, 10, None, None, 0, None)
    assert rc == -1 and b"null" in lib.vn_last_error()
    rc = lib.vn_vfe_layer_fwd(None, None, 5, 35, 9, 16, None, None, None, None, None, None)
    assert rc < 0
    p16 = ctypes.c_void_p(16)
    rc = lib.vn_conv3d_bn_relu_fwd(p16, 1, 10, 16, 12, 128, 2, 1, p16, p16, p16, 0, None, 0, p16, None, 0, None, p16, 1 << 20, None)
    assert rc == -1 and b"multiple of 8" in lib.vn_last_error()
    assert lib.vn_conv3d_bn_relu_fwd(p16, 1, 10, 16, 16, 128, 2, 1, p16, p16, p16, 0, None, 0, p16, None, 0, None, p16, 8, None) == -2
    assert lib.vn_conv3d_out_state_bytes(16, 400, 352) == 16 * 25 * 44
    assert lib.vn_conv3d_workspace_bytes(16, 400, 352) == 16 + 16 * 25 * 44 * (16 + 8 + 8 * 8)
    assert lib.vn_conv3d_box_flags_bytes(16, 10, 400, 352) == 16 * 10 * 25 * 44 * 3
    assert lib.vn_conv3d_mark_boxes(None, 5, None, 1, 10, 16, 16, 4, None, None, None, None) == -1
    assert lib.vn_conv3d_packed_weight_bytes(128) == 27 * 128 * 64 * 2 and lib.vn_conv3d_packed_weight_bytes(32) == 0
    assert lib.vn_bbox_iou(None, None, 0, 5, None, None) == 0  # empty: no-op
    assert lib.vn_bbox_iou(None, p16, 3, 5, p16, None) == -1 and b"null" in lib.vn_last_error()
    assert lib.vn_bbox_iou(ctypes.c_void_p(20), p16, 3, 5, p16, None) == -1 and b"aligned" in lib.vn_last_error()
    need = lib.vn_rpn_targets_workspace_bytes(2, 70400, 9)
    assert need >= 9 * 8 + 2 * 70400 * 8 and lib.vn_rpn_targets_workspace_bytes(-1, 4, 4) == 0
    targ = (p16, p16, 70400, p16, p16, p16, 2, 9, 0.6, 0.45, 1.56, p16, p16, p16, p16)
    assert lib.vn_rpn_targets(*targ, need - 1, None) == -1 and b"workspace" in lib.vn_last_error()
    assert lib.vn_rpn_targets(p16, p16, 70400, None, None, p16, 2, 9, 0.6, 0.45, 1.56, p16, p16, p16, p16, need, None) == -1
    with pytest.raises(_cabi.VnError):
        _cabi.check(rc, "probe")


def test_product_path_fails_loudly_without_cuda():
    import voxelnet_pytorch_b200 as vn

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError, match="CUDA"):
        vn.pcl_to_voxels(np.zeros((10, 4), np.float32), "Car")
    net = vn.FeatureLearningNet().eval()
    with pytest.raises(RuntimeError, match="CUDA"):
        net.voxelwise(torch.zeros((3, 35, 7)))
    with pytest.raises(RuntimeError, match="CUDA"):
        vn.MiddleConvNet().eval()(torch.zeros((1, 10, 16, 8, 128)))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "voxelnet-pytorch_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("# oracle", ""), "%s mentions the oracle" % f


def test_prepare_voxel_and_collate():
    import voxelnet_pytorch_b200 as vn
    from oracle import voxelize_np as OV
    from tests import golden_util as G

    pts = G.kitti_sub_points()
    d = [OV.pcl_to_voxels_fast(pts[:4000], "Car"), OV.pcl_to_voxels_fast(pts[4000:9000], "Car"), OV.pcl_to_voxels_fast(pts[:0], "Car")]
    f, n, c = vn.prepare_voxel(d)
    f0, n0, c0 = OV.prepare_voxel(d)
    for a, b in zip(c, c0):
        assert a.dtype == np.int64 and np.array_equal(a, b)
    assert f[1] is d[1]["feature_buffer"] and c[2].shape == (0, 4)
    parts = [("t%d" % i, None, pts[:10], [], d[i]) for i in range(3)]
    out = vn.collate_fn(parts)
    assert len(out) == 7 and out[2][0].dtype == torch.float32 and out[4][1].dtype == torch.int64
    assert torch.equal(out[4][1], torch.from_numpy(c0[1]))


def test_modules_keep_reference_parameter_names_and_shapes():
    import voxelnet_pytorch_b200 as vn
    from oracle import model_torch as OM

    sd = OM.random_state_dict(0)
    f, m = vn.FeatureLearningNet(), vn.MiddleConvNet()
    f.load_state_dict({k: v for k, v in sd.items() if k.startswith("vfe_")})
    m.load_state_dict({k: v for k, v in sd.items() if k.startswith("middle_layer.")})
    assert f.vfe_1.fcn[0].weight.shape == (16, 7) and f.vfe_2.fcn[0].weight.shape == (64, 32)
    assert m.middle_layer[0].conv.weight.shape == (64, 128, 3, 3, 3)
    assert m.middle_layer[1].conv.stride == (1, 1, 1) and m.middle_layer[1].conv.padding == (0, 1, 1)
    with pytest.raises(ValueError):
        vn.ConvMD(2, 128, 128, 3, (1, 1), (1, 1))
    with pytest.raises(RuntimeError):  # train mode runs the statistics kernel: CUDA only, like every entry point
        vn.VFELayer(7, 32).train()(torch.zeros(1, 35, 7), torch.ones(1, 35, 1, dtype=torch.bool))


def test_bn_folding_is_the_eval_affine():
    import voxelnet_pytorch_b200 as vn
    from oracle import model_torch as OM

    sd = OM.random_state_dict(4)
    l = vn.VFELayer(7, 32).eval()
    l.load_state_dict({k[len("vfe_1."):]: v for k, v in sd.items() if k.startswith("vfe_1.")})
    w, b, s, t = l.folded()
    x = torch.randn(50, 7)
    want = torch.nn.functional.batch_norm(torch.relu(x @ w.T + b), sd["vfe_1.bn.running_mean"], sd["vfe_1.bn.running_var"],
                                          sd["vfe_1.bn.weight"], sd["vfe_1.bn.bias"], False, 0.0, 1e-5)
    assert torch.allclose(torch.relu(x @ w.T + b) * s + t, want, rtol=1e-5, atol=1e-6)


def test_synthetic_frames_are_kitti_shaped():
    from oracle import voxelize_np as OV
    from voxelnet_pytorch_b200 import synth

    p = synth.lidar_frame(0)
    assert p.dtype == np.float32 and p.shape[1] == 4 and 100_000 < len(p) < 135_000
    assert np.array_equal(p, synth.lidar_frame(0))  # seeded
    v = OV.pcl_to_voxels_fast(p, "Car")
    K = len(v["number_buffer"])
    assert 9_000 < K < 16_000, K  # KITTI-typical 12-14k
    assert 3.5 < v["number_buffer"].mean() < 6.5


def test_raw_point_input_pipeline_host_side(tmp_path):
    """dataset.py:51 loader and the pinned CSR packing of the raw-point input pipeline."""
    import numpy as np

    from voxelnet_pytorch_b200 import PinnedFrameBatch, load_velodyne_bin

    rng = np.random.default_rng(3)
    frames = [rng.standard_normal((n, 4)).astype(np.float32) for n in (5, 0, 17)]
    paths = []
    for i, f in enumerate(frames):
        p = tmp_path / ("%06d.bin" % i)
        f.tofile(p)
        paths.append(str(p))
    assert np.array_equal(load_velodyne_bin(paths[2]), frames[2])
    hb = PinnedFrameBatch.from_files(paths)
    assert hb.n_points == 22 and hb.max_points_per_frame == 17
    assert hb.offsets.tolist() == [0, 5, 5, 22]
    assert np.array_equal(hb.points.numpy()[:22], np.concatenate(frames))
    hb.fill([frames[2][:3], frames[0], frames[1]])  # reuse of the pinned allocation
    assert hb.offsets.tolist() == [0, 3, 8, 8] and hb.n_points == 8
    with pytest.raises(ValueError):
        hb.fill(frames[:2])
    with pytest.raises(ValueError):
        hb.fill([np.zeros((30, 4), np.float32)] * 3)
    (tmp_path / "bad.bin").write_bytes(b"\0" * 20)
    with pytest.raises(ValueError):
        load_velodyne_bin(str(tmp_path / "bad.bin"))


def test_host_frame_packer_packs_ragged_frames_and_is_reusable():
    """PinnedFrameBatch.fill / fill_async (vn_host_pack_submit / _wait, the library's host thread pool): ragged and
    empty frames, non-float32 / non-contiguous inputs, reuse of the same buffer, capacity and shape errors."""
    import voxelnet_pytorch_b200 as vn

    rng = np.random.default_rng(3)
    frames = [rng.random((n, 4), dtype=np.float32) for n in (70001, 0, 5, 333333, 1)]
    pb = vn.PinnedFrameBatch(5, 500000)
    pb.fill(frames)
    want = np.concatenate(frames, 0)
    assert pb.n_points == len(want) and pb.max_points_per_frame == 333333
    assert pb.offsets.tolist() == [0, 70001, 70001, 70006, 403339, 403340]
    assert np.array_equal(pb.points.numpy()[: pb.n_points], want)
    # reuse with other dtypes / strides, asynchronously, several jobs in flight
    others = [vn.PinnedFrameBatch(2, 1000) for _ in range(3)]
    srcs = [[rng.random((300, 8))[:, ::2], rng.random((17, 4)).astype(np.float64)] for _ in others]
    for o, fr in zip(others, srcs):
        o.fill_async(fr)
    pb.fill_async(frames[::-1])
    for o, fr in zip(others, srcs):
        o.wait_filled()
        assert np.array_equal(o.points.numpy()[: o.n_points], np.concatenate([f.astype(np.float32) for f in fr], 0))
    pb.wait_filled()
    assert np.array_equal(pb.points.numpy()[: pb.n_points], np.concatenate(frames[::-1], 0))
    with pytest.raises(ValueError):
        pb.fill(frames[:4])
    with pytest.raises(ValueError):
        vn.PinnedFrameBatch(1, 10).fill([rng.random((11, 4), dtype=np.float32)])
    with pytest.raises(ValueError):
        vn.PinnedFrameBatch(1, 10).fill([rng.random((5, 3), dtype=np.float32)])
    empty = vn.PinnedFrameBatch(2, 4).fill([np.zeros((0, 4), np.float32)] * 2)
    assert empty.n_points == 0 and empty.offsets.tolist() == [0, 0, 0]


def test_load_velodyne_bin_pinned_reads_the_same_points():
    """dataset.load_velodyne_bin(pinned=True) (page-locked on a CUDA host, plain memory here) against np.fromfile."""
    import os

    import voxelnet_pytorch_b200 as vn
    from tests import golden_util as GU

    path = os.path.join(GU.GOLDEN, "kitti_0000000000.bin")
    a = vn.load_velodyne_bin(path, pinned=True)
    b = vn.load_velodyne_bin(path)
    assert a.dtype == np.float32 and a.shape == b.shape == (os.path.getsize(path) // 16, 4)
    assert np.array_equal(a, b)
    assert vn.pinned_frame(7).shape == (7, 4)
