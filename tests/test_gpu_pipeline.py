"""GPU parity: the fused front end (points -> (B,128,H,W)) vs the oracle chain, and state reuse."""
import numpy as np
import pytest
import torch

from oracle import model_torch as OM
from tests.helpers import load_state, oracle_batch, rel_max_err, synth_frames

pytestmark = pytest.mark.gpu


def _build(cuda_device, target, sd, **kw):
    import voxelnet_pytorch_b200 as vn

    fnet = vn.FeatureLearningNet(target).eval()
    mnet = vn.MiddleConvNet().eval()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    return vn.VoxelNetFrontEnd(fnet.to(cuda_device), mnet.to(cuda_device), target, **kw)


def test_front_end_matches_oracle_ped(cuda_device):
    """Full chain on the Pedestrian grid (smaller: the CPU oracle conv finishes in seconds)."""
    sd = OM.random_state_dict(11)
    frames = synth_frames(2, 30)
    fe = _build(cuda_device, "Pedestrian", sd)
    got = fe.forward_frames(frames).cpu()
    offs = fe.voxel_counts()
    _, feats, _, coords = oracle_batch(frames, "Pedestrian")
    assert list(np.diff(offs)) == [len(f) for f in feats]
    dense = OM.feature_learning_net([torch.from_numpy(f) for f in feats], [torch.from_numpy(c) for c in coords], sd, "Pedestrian")
    want = OM.middle_layers(dense, sd)
    assert got.shape == want.shape == (2, 128, 200, 240)
    assert rel_max_err(got, want) < 1e-2
    # second call with different frames reuses the resident grid: it must have been re-zeroed
    frames2 = synth_frames(2, 40)
    got2 = fe.forward_frames(frames2).cpu()
    _, feats2, _, coords2 = oracle_batch(frames2, "Pedestrian")
    dense2 = OM.feature_learning_net([torch.from_numpy(f) for f in feats2], [torch.from_numpy(c) for c in coords2], sd, "Pedestrian")
    want2 = OM.middle_layers(dense2, sd)
    assert rel_max_err(got2, want2) < 1e-2
    # default policy: a grid this small is materialised (resident, re-zeroed by the step itself) ...
    assert fe.materialise_grid and fe.rows is None and int(fe.dense.view(torch.int16).ne(0).sum()) == 0
    # ... the never-materialised form (row table + occupancy index) gives the same bits
    fe_rows = _build(cuda_device, "Pedestrian", sd, materialise_grid=False)
    assert torch.equal(fe_rows.forward_frames(frames).cpu(), got)
    assert torch.equal(fe_rows.forward_frames(frames2).cpu(), got2)
    assert fe_rows.dense is None and fe_rows.rows is not None
    # and it is what the default policy picks when the grid would not fit the budget
    fe_auto = _build(cuda_device, "Pedestrian", sd, grid_budget_bytes=1 << 20)
    assert torch.equal(fe_auto.forward_frames(frames2).cpu(), got2) and fe_auto.dense is None


def test_front_end_matches_oracle_car(cuda_device):
    """Full chain on the Car grid (two full-size frames) against the oracle, plus determinism.  The benchmarked
    configuration (4-frame batches, graph replay, three front ends in flight) is in tests/test_gpu_bench_parity.py."""
    sd = OM.random_state_dict(12)
    frames = synth_frames(2, 50)
    fe = _build(cuda_device, "Car", sd)
    a = fe.forward_frames(frames).clone()
    b = fe.forward_frames(frames)
    assert a.shape == (2, 128, 400, 352)
    assert torch.equal(a, b)
    _, feats, _, coords = oracle_batch(frames, "Car")
    dense = OM.feature_learning_net([torch.from_numpy(f) for f in feats], [torch.from_numpy(c) for c in coords], sd, "Car")
    want = OM.middle_layers(dense, sd)
    assert rel_max_err(a.cpu(), want) < 1e-2


@pytest.mark.parametrize("target", ["Car", "Pedestrian"])
def test_fused_voxel_svfe_is_bit_identical_to_materialised_path(cuda_device, target):
    """vn_voxelize_svfe_fused (no feature buffer) == vn_voxelize_fill + vn_svfe_fused_fwd, bit for bit,
    and both match the oracle; coordinates / counts identical."""
    import voxelnet_pytorch_b200 as vn
    from voxelnet_pytorch_b200 import _cabi, ops

    sd = OM.random_state_dict(13)
    frames = synth_frames(3, 60)
    frames[1] = frames[1][:5000]
    # a voxel with > T points and one with > 32 points exercises both selection paths
    rng = np.random.default_rng(0)
    blob = np.stack([rng.uniform(20.0, 20.19, 300), rng.uniform(1.0, 1.19, 300), rng.uniform(-1.0, -0.61, 300), rng.random(300)], 1)
    frames[2] = np.concatenate([frames[2], blob.astype(np.float32)])
    fnet = vn.FeatureLearningNet(target).eval()
    load_state(fnet, sd, "")
    fnet = fnet.to(cuda_device)
    points, offsets, max_n = vn.pack_frames(frames)
    vb = vn.voxelize_batch(points, offsets, max_n, target)
    want = fnet.voxelwise(vb.feature)
    grid = _cabi.grid_for_target(target)
    ws = torch.empty(ops.voxelize_workspace_bytes(grid, 3, points.shape[0]), dtype=torch.uint8, device=cuda_device)
    vo = torch.empty(4, dtype=torch.int32, device=cuda_device)
    ops.voxelize_index(grid, points, offsets, max_n, vo, ws)
    K = vb.feature.shape[0]
    vw = torch.empty((K, 128), dtype=torch.float32, device=cuda_device)
    co = torch.empty((K, 4), dtype=torch.int64, device=cuda_device)
    nu = torch.empty((K,), dtype=torch.int64, device=cuda_device)
    dense = torch.zeros((3, grid.D, grid.H, grid.W, 128), dtype=torch.bfloat16, device=cuda_device)
    ops.voxelize_svfe_fused(grid, points, offsets, vo, fnet.vfe_1.folded(), fnet.vfe_2.folded(), vw, co, nu, ws, dense=dense)
    # the kernel's own scatter == vn_scatter_dense_bf16 of its voxelwise rows
    dense2 = torch.zeros_like(dense)
    ops.scatter_dense_bf16(want, vb.coordinate, dense2)
    assert torch.equal(dense.view(torch.int16), dense2.view(torch.int16))
    # dense-only mode (no voxelwise buffer) writes the same cells
    dense3 = torch.zeros_like(dense)
    co3 = torch.empty_like(co)
    ops.voxelize_svfe_fused(grid, points, offsets, vo, fnet.vfe_1.folded(), fnet.vfe_2.folded(), None, co3, None, ws, dense=dense3)
    assert torch.equal(dense3.view(torch.int16), dense2.view(torch.int16)) and torch.equal(co3, co)
    assert torch.equal(vo.cpu(), vb.voxel_offsets.cpu())
    assert torch.equal(co, vb.coordinate) and torch.equal(nu, vb.number)
    assert torch.equal(vw.view(torch.int32), want.view(torch.int32))
    _, feats, _, _ = oracle_batch(frames, target)
    ref = OM.svfe_voxelwise(torch.from_numpy(np.concatenate(feats)), sd)
    assert torch.allclose(vw.cpu(), ref, rtol=1e-4, atol=1e-5)


def test_front_end_fused_and_unfused_agree(cuda_device):
    sd = OM.random_state_dict(14)
    frames = synth_frames(2, 70)
    a = _build(cuda_device, "Pedestrian", sd, fuse_voxel_features=True).forward_frames(frames).clone()
    b = _build(cuda_device, "Pedestrian", sd, fuse_voxel_features=False).forward_frames(frames)
    assert torch.equal(a, b)


def test_front_end_sparse_and_dense_chains_agree(cuda_device):
    """skip_empty_boxes=True (live-box skipping + background deltas in all three layers) vs the dense
    plain chain on real voxelized frames: equal up to the bf16 rounding of the two intermediate
    activations; the live-box tables are consumed (all zero) after the call."""
    sd = OM.random_state_dict(15)
    frames = synth_frames(2, 80)
    fe = _build(cuda_device, "Pedestrian", sd, skip_empty_boxes=True)
    a = fe.forward_frames(frames).clone()
    assert all(int(f.sum()) == 0 for f in fe.box_flags)
    b = _build(cuda_device, "Pedestrian", sd, skip_empty_boxes=False).forward_frames(frames)
    assert rel_max_err(a, b) < 6e-3
    assert float((a - b).norm() / b.norm()) < 3e-3


@pytest.mark.parametrize("target", ["Pedestrian", "Car"])
def test_resident_outputs_are_bit_identical_over_a_frame_sequence(cuda_device, target):
    """resident_outputs=True stores only the output blocks a voxel reaches now or reached in the previous
    step; over a sequence of different batches (dense, nearly empty, dense again) every result must
    equal, bit for bit, the result of the pipeline that rewrites every block."""
    sd = OM.random_state_dict(16)
    a, b = synth_frames(2, 90), synth_frames(2, 95)
    sparse = [f[:300].copy() for f in b]
    seq = [a, b, sparse, a, a]
    res = _build(cuda_device, target, sd, resident_outputs=True)
    full = _build(cuda_device, target, sd, resident_outputs=False)
    for i, frames in enumerate(seq):
        got = res.forward_frames(frames)
        want = full.forward_frames(frames)
        assert torch.equal(got.view(torch.int32), want.view(torch.int32)), "step %d" % i
        assert torch.equal(res.act1.view(torch.int16), full.act1.view(torch.int16)), "act1 step %d" % i
        assert torch.equal(res.act2.view(torch.int16), full.act2.view(torch.int16)), "act2 step %d" % i
    stats = res.plan_stats()
    assert all(s["restored_blocks"] == 0 for s in stats)  # the last two steps saw the same frames
    assert all(0 < s["touched_blocks"] < s["blocks_total"] for s in stats)
    # new parameters change the constants the untouched blocks hold: refresh must invalidate them
    sd2 = OM.random_state_dict(17)
    for fe in (res, full):
        load_state(fe.feature_net, sd2, "")
        load_state(fe.middle_net, sd2, "")
        fe.refresh_parameters()
    assert torch.equal(res.forward_frames(b), full.forward_frames(b))


def test_frames_are_independent_units_including_an_empty_one(cuda_device):
    """Size-independent property at full Car size: a batch [A, B, empty, C] gives, frame by frame and bit
    for bit, what each frame gives alone (no stage couples frames: eval BatchNorm, per-frame voxelization,
    no cross-batch term in the convolutions) - the property the multi-GPU sharding rests on.  The empty
    frame yields the all-background output, equal to a frame whose points are all outside the crop."""
    sd = OM.random_state_dict(18)
    a, b, c = synth_frames(3, 120)
    empty = np.zeros((0, 4), np.float32)
    far = np.array([[500.0, 0.0, 0.0, 0.5], [np.nan, 1.0, 1.0, 0.1], [10.0, np.inf, 0.0, 0.2]], np.float32)
    fe = _build(cuda_device, "Car", sd)
    batch = fe.forward_frames([a, b, empty, c]).clone()
    assert list(np.diff(fe.voxel_counts()))[2] == 0
    single = _build(cuda_device, "Car", sd)
    for i, f in enumerate([a, b, far, c]):
        alone = single.forward_frames([f])
        assert torch.equal(alone[0].view(torch.int32), batch[i].view(torch.int32)), "frame %d" % i
    assert bool(torch.isfinite(batch).all())


def test_host_frame_stream_matches_direct_forward(cuda_device):
    """The end-to-end feeding API (pinned batches, two staging slots, copy stream) returns what forward_points
    returns on the same frames, batch after batch, and reports the per-frame voxel counts."""
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(9)
    fnet, mnet = vn.FeatureLearningNet("Pedestrian").eval(), vn.MiddleConvNet().eval()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    fe = vn.VoxelNetFrontEnd(fnet.to(cuda_device), mnet.to(cuda_device), "Pedestrian")
    sets = [[f[::5].copy() for f in synth_frames(2, s)] for s in (300, 310, 320)]
    want, counts = [], []
    for fr in sets:
        want.append(fe.forward_frames(fr).clone())
        counts.append(fe.voxel_counts())
    hbs = [vn.PinnedFrameBatch.from_frames(fr) for fr in sets]
    stream = vn.HostFrameStream(fe, 2, max(hb.n_points for hb in hbs))
    stream.upload(hbs[0])
    for i in range(3):
        if i + 1 < 3:
            stream.upload(hbs[i + 1])
        out = stream.forward()
        got_counts = stream.voxel_counts_host().tolist()
        assert got_counts == counts[i]
        assert torch.equal(out, want[i]), i
    with pytest.raises(RuntimeError):
        stream.forward()
    # the same batches as per-frame arrays: page-locked ones are copied frame by frame with no host packing
    # (vn_upload_frames), pageable ones go through the library's packer - same tensors, same counts
    import os

    from tests import golden_util as GU

    kitti = vn.load_velodyne_bin(os.path.join(GU.GOLDEN, "kitti_0000000000.bin"), pinned=True)
    assert torch.from_numpy(kitti).is_pinned()
    assert np.array_equal(kitti, np.fromfile(os.path.join(GU.GOLDEN, "kitti_0000000000.bin"), dtype=np.float32).reshape(-1, 4))
    for pinned in (True, False):
        def host(f):
            if not pinned:
                return f.copy()
            a = vn.pinned_frame(f.shape[0])
            a[...] = f
            return a

        assert torch.from_numpy(host(sets[0][0])).is_pinned() == pinned
        stream.upload_frames([host(f) for f in sets[0]])
        for i in range(3):
            if i + 1 < 3:
                stream.upload_frames([host(f) for f in sets[i + 1]])
            out = stream.forward()
            assert stream.voxel_counts_host().tolist() == counts[i]
            assert torch.equal(out, want[i]), (pinned, i)
    with pytest.raises(ValueError):
        stream.upload_frames([sets[0][0]])  # wrong frame count


def test_cuda_graph_replay_equals_eager(cuda_device):
    """capture_graph / HostFrameStream(cuda_graphs=True): replays of the captured step give bit-identical tensors
    and voxel counts to eager launches, with graphs of different batches interleaved."""
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(10)
    fnet, mnet = vn.FeatureLearningNet("Pedestrian").eval(), vn.MiddleConvNet().eval()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    fe = vn.VoxelNetFrontEnd(fnet.to(cuda_device), mnet.to(cuda_device), "Pedestrian")
    sets = [[f[::5].copy() for f in synth_frames(2, s)] for s in (400, 410)]
    packed = [vn.pack_frames(fr) for fr in sets]
    packed.sort(key=lambda p: -p[0].shape[0])
    want = [fe.forward_points(*p).clone() for p in packed]
    graphs = [fe.capture_graph(*p) for p in packed]
    for i in (0, 1, 1, 0, 1):
        graphs[i].replay()
        torch.cuda.synchronize()
        assert torch.equal(fe.out, want[i]), i
    hbs = [vn.PinnedFrameBatch.from_frames(fr) for fr in sets]
    stream = vn.HostFrameStream(fe, 2, max(hb.n_points for hb in hbs), cuda_graphs=True)
    ref = [fe.forward_frames(fr).clone() for fr in sets]
    stream.upload(hbs[0])
    for i in range(6):  # slots alternate: each (slot, batch) pair is seen eagerly once, then captured, then replayed
        if i + 1 < 6:
            stream.upload(hbs[(i + 1) % 2])
        out = stream.forward()
        stream.voxel_counts_host()
        assert torch.equal(out, ref[i % 2]), i
    assert len(stream._graphs) == 2 and not torch.equal(ref[0], ref[1])


def test_host_stream_lagged_counts(cuda_device):
    """voxel_counts_host(lag=1) returns the previous step's counts without waiting for the running step."""
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(11)
    fnet, mnet = vn.FeatureLearningNet("Pedestrian").eval(), vn.MiddleConvNet().eval()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    fe = vn.VoxelNetFrontEnd(fnet.to(cuda_device), mnet.to(cuda_device), "Pedestrian")
    sets = [[f[::6].copy() for f in synth_frames(2, s)] for s in (500, 510, 520)]
    want = []
    for fr in sets:
        fe.forward_frames(fr)
        want.append(fe.voxel_counts())
    hbs = [vn.PinnedFrameBatch.from_frames(fr) for fr in sets]
    stream = vn.HostFrameStream(fe, 2, max(hb.n_points for hb in hbs))
    with pytest.raises(RuntimeError):
        stream.voxel_counts_host()
    stream.upload(hbs[0])
    got = []
    for i in range(3):
        if i + 1 < 3:
            stream.upload(hbs[i + 1])
        stream.forward()
        if i == 0:
            with pytest.raises(RuntimeError):
                stream.voxel_counts_host(lag=1)
        else:
            got.append(stream.voxel_counts_host(lag=1).tolist())
    got.append(stream.voxel_counts_host().tolist())
    assert got == want


@pytest.mark.parametrize("graphs", [False, True])
def test_host_stream_with_two_front_ends_in_flight(cuda_device, graphs):
    """HostFrameStream over TWO front ends (two steps in flight on two streams, three staging slots): every step's
    tensor - consumed on the caller's stream before the next call - and its voxel counts (read up to two steps
    behind) are bit-identical to the single-front-end run of the same batch sequence."""
    import voxelnet_pytorch_b200 as vn

    sd = OM.random_state_dict(12)
    fnet, mnet = vn.FeatureLearningNet("Pedestrian").eval(), vn.MiddleConvNet().eval()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    fnet, mnet = fnet.to(cuda_device), mnet.to(cuda_device)
    fes = [vn.VoxelNetFrontEnd(fnet, mnet, "Pedestrian") for _ in range(3)]
    sets = [[f[::5].copy() for f in synth_frames(2, s)] for s in (600, 610, 620)]
    want, counts = [], []
    for fr in sets:
        want.append(fes[2].forward_frames(fr).clone())
        counts.append(fes[2].voxel_counts())
    assert not torch.equal(want[0], want[1])
    hbs = [vn.PinnedFrameBatch.from_frames(fr) for fr in sets]
    stream = vn.HostFrameStream(fes[:2], 2, max(hb.n_points for hb in hbs), cuda_graphs=graphs)
    order = [0, 1, 2, 1, 0, 2, 2, 1] * (4 if graphs else 1)  # with graphs: long enough for captures and replays
    n = len(order)
    stream.upload(hbs[order[0]])
    stream.upload(hbs[order[1]])
    checksum = torch.zeros((), dtype=torch.float64, device=cuda_device)
    got_counts = []
    for i in range(n):
        if i + 2 < n:
            stream.upload(hbs[order[i + 2]])
        out = stream.forward()
        assert torch.equal(out, want[order[i]]), i                 # consumed on the current stream, before the next call
        checksum += out.double().sum()                             # (and an asynchronous consumer as well)
        if i >= 2:
            got_counts.append(stream.voxel_counts_host(lag=2).tolist())
    got_counts += [stream.voxel_counts_host(lag=1).tolist(), stream.voxel_counts_host().tolist()]
    assert got_counts == [counts[b] for b in order]
    assert float(checksum) == float(sum(want[b].double().sum() for b in order))
    if graphs:
        assert len(stream._graphs) > 0
    with pytest.raises(RuntimeError):
        stream.voxel_counts_host(lag=3)


def test_bf16_channels_last_output_option(cuda_device):
    """output_format="nhwc_bf16": the same numbers as the float32 drop-in tensor, rounded to bf16, in channels-last memory
    with the documented channel permutation (what a bf16 channels-last RPN consumes without a cast / layout pass);
    resident outputs over a frame sequence included."""
    sd = OM.random_state_dict(21)
    sets = [synth_frames(2, s) for s in (500, 510, 500)]
    fe32 = _build(cuda_device, "Pedestrian", sd)
    fe16 = _build(cuda_device, "Pedestrian", sd, output_format="nhwc_bf16")
    perm = fe16.reference_channel_of_output().to(cuda_device)
    assert torch.equal(fe32.reference_channel_of_output(), torch.arange(128))
    for fr in sets:
        a = fe32.forward_frames(fr)
        b = fe16.forward_frames(fr)
        assert b.dtype == torch.bfloat16 and b.shape == a.shape and b.is_contiguous(memory_format=torch.channels_last)
        assert torch.equal(b, a[:, perm].to(torch.bfloat16))
    # a convolution with its input channels permuted once sees the same tensor
    w = torch.randn(8, 128, 3, 3, device=cuda_device)
    ref = torch.nn.functional.conv2d(a.to(torch.bfloat16).float(), w, padding=1)
    got = torch.nn.functional.conv2d(b.float(), w[:, perm], padding=1)
    assert rel_max_err(got.cpu(), ref.cpu()) < 1e-5
