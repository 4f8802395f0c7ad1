"""The stride-1 3x3 RPN layers (reference ConvMD(2, 128, 128, 3, (1,1), (1,1)), voxelnet/model.py:111-167, 212-242) on the
tcgen05 conv kernel: `vn_conv2d_bn_relu_fwd` / `ConvMD2d` / `RPNBlockTail` against plain PyTorch fp32.

Tolerances: bf16 operands, fp32 accumulation, bf16 output.  Against the SAME bf16-rounded operands the result must be
the fp32 result rounded to bf16 up to one bf16 ulp (catches any wrong tap, swizzle, channel group or descriptor);
against the unrounded fp32 module (Conv2d -> BatchNorm2d -> ReLU) max|err| / max|ref| <= 1e-2 like the 3-D layers.
"""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from tests.helpers import rel_max_err

pytestmark = pytest.mark.gpu


class RefConvMD2d(nn.Module):
    """What the reference's ConvMD(2, cin, cout, k, stride, padding) is (model.py:132-167)."""

    def __init__(self, cin, cout, k=3, stride=1, padding=1):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, k, stride, padding)
        self.batch_norm = nn.BatchNorm2d(cout)

    def forward(self, x):
        return F.relu(self.batch_norm(self.conv(x)))


def _randomise(m, g):
    for mod in m.modules():
        if isinstance(mod, nn.BatchNorm2d):
            mod.weight.data = 0.5 + torch.rand(mod.weight.shape, generator=g)
            mod.bias.data = (torch.rand(mod.bias.shape, generator=g) - 0.5) * 0.4
            mod.running_mean.data = (torch.rand(mod.running_mean.shape, generator=g) - 0.5) * 0.3
            mod.running_var.data = 0.5 + torch.rand(mod.running_var.shape, generator=g)
    return m.eval()


@pytest.mark.parametrize("cin,cout,B,H,W", [(128, 128, 2, 24, 16), (64, 64, 1, 16, 8), (128, 192, 1, 37, 24), (64, 128, 3, 100, 88),
                                              (256, 256, 2, 50, 44), (128, 64, 1, 9, 13)])
def test_conv2d_bn_relu_matches_torch(cuda_device, cin, cout, B, H, W):
    import voxelnet_pytorch_b200 as vn

    g = torch.Generator().manual_seed(cin + cout + H)
    ref = _randomise(RefConvMD2d(cin, cout), g)
    x = (torch.randn((B, cin, H, W), generator=g) * (torch.rand((B, cin, H, W), generator=g) < 0.6)).to(torch.bfloat16)
    layer = vn.ConvMD2d.from_reference(ref.to(cuda_device))
    assert set(layer.state_dict()) == set(ref.state_dict())
    xg = x.to(cuda_device).contiguous(memory_format=torch.channels_last)
    got = layer(xg)
    assert got.dtype == torch.bfloat16 and got.shape == (B, cout, H, W) and got.is_contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        want = ref.cpu().double()(x.double())
        # the same arithmetic on the operands the kernel sees: weights * BN scale rounded to bf16, fp32-exact accumulation
        bn = ref.batch_norm
        scale = (bn.weight.double() / torch.sqrt(bn.running_var.double() + bn.eps)).float()
        w_q = (ref.conv.weight.float() * scale[:, None, None, None]).to(torch.bfloat16).double()
        shift = (ref.conv.bias.double() * scale.double() + bn.bias.double() - bn.running_mean.double() * scale.double()).float().double()
        exact = F.relu(F.conv2d(x.double(), w_q, None, 1, 1) + shift[None, :, None, None])
    got64 = got.cpu().double()
    assert rel_max_err(got64, want) < 1e-2
    ulp = exact.abs().clamp_min(2.0 ** -100) * 2.0 ** -7  # one bf16 ulp is at most 2^-7 of the value
    # (+ 2e-5: fp32 accumulation over up to 2,304 products before the rounding to bf16)
    assert bool(((got64 - exact).abs() <= ulp + 2e-5).all()), float(((got64 - exact).abs() - ulp).max())


@pytest.mark.parametrize("cin,cout,B,H,W", [(128, 128, 2, 32, 16), (128, 128, 1, 50, 44), (64, 64, 1, 17, 9), (128, 256, 2, 40, 48)])
def test_conv2d_stride2_matches_torch(cuda_device, cin, cout, B, H, W):
    """The heads of the RPN blocks: ConvMD(2, 128, 128|256, 3, (2,2), (1,1)) (model.py:213,233,244).  Odd / even input rows
    travel as two boxes; odd input extents give (H-1)/2+1 outputs like torch."""
    import voxelnet_pytorch_b200 as vn

    g = torch.Generator().manual_seed(cin + cout + H + 1)
    ref = _randomise(RefConvMD2d(cin, cout, 3, 2, 1), g)
    x = (torch.randn((B, cin, H, W), generator=g) * (torch.rand((B, cin, H, W), generator=g) < 0.6)).to(torch.bfloat16)
    layer = vn.ConvMD2d.from_reference(ref.to(cuda_device))
    got = layer(x.to(cuda_device).contiguous(memory_format=torch.channels_last))
    with torch.no_grad():
        want = ref.cpu().double()(x.double())
        bn = ref.batch_norm
        scale = (bn.weight.double() / torch.sqrt(bn.running_var.double() + bn.eps)).float()
        w_q = (ref.conv.weight.float() * scale[:, None, None, None]).to(torch.bfloat16).double()
        shift = (ref.conv.bias.double() * scale.double() + bn.bias.double() - bn.running_mean.double() * scale.double()).float().double()
        exact = F.relu(F.conv2d(x.double(), w_q, None, 2, 1) + shift[None, :, None, None])
    assert got.shape == want.shape and got.dtype == torch.bfloat16
    got64 = got.cpu().double()
    assert rel_max_err(got64, want) < 1e-2
    ulp = exact.abs().clamp_min(2.0 ** -100) * 2.0 ** -7
    assert bool(((got64 - exact).abs() <= ulp + 2e-5).all()), float(((got64 - exact).abs() - ulp).max())


@pytest.mark.parametrize("cin,cout,H,W", [(128, 256, 40, 32), (64, 192, 20, 16)])
def test_deconv_3x3_stride1_matches_torch(cuda_device, cin, cout, H, W):
    """The reference's deconv1 (DeConv2d(128, 256, 3, (1,1), (1,1)), model.py:229): ConvTranspose2d -> BN -> ReLU, 256
    output channels = two launches of 128 into one tensor."""
    import voxelnet_pytorch_b200 as vn

    class RefDeConv(nn.Module):
        def __init__(self):
            super().__init__()
            self.deconv = nn.ConvTranspose2d(cin, cout, 3, 1, 1)
            self.batch_norm = nn.BatchNorm2d(cout)

        def forward(self, x):
            return F.relu(self.batch_norm(self.deconv(x)))

    g = torch.Generator().manual_seed(cout)
    ref = _randomise(RefDeConv(), g)
    x = torch.randn((2, cin, H, W), generator=g).relu().to(torch.bfloat16)
    layer = vn.DeConvMD2d.from_reference(ref.to(cuda_device))
    assert set(layer.state_dict()) == set(ref.state_dict())
    got = layer(x.to(cuda_device).contiguous(memory_format=torch.channels_last))
    with torch.no_grad():
        want = ref.cpu().double()(x.double())
    assert got.shape == want.shape and got.dtype == torch.bfloat16
    assert rel_max_err(got.cpu().double(), want) < 1e-2
    assert float((got.cpu().double() - want).norm() / want.norm()) < 5e-3


@pytest.mark.parametrize("cin,cout,k,H,W", [(128, 256, 2, 20, 24), (256, 256, 4, 13, 11), (64, 64, 3, 8, 8)])
def test_deconv_kernel_equals_stride_matches_torch(cuda_device, cin, cout, k, H, W):
    """deconv2 / deconv3 (DeConv2d(128, 256, 2, (2,2), (0,0)), DeConv2d(256, 256, 4, (4,4), (0,0)), model.py:242,253) as k*k
    pointwise launches that write each output pixel once, into a channel slice of a wider tensor."""
    import voxelnet_pytorch_b200 as vn

    class RefDeConv(nn.Module):
        def __init__(self):
            super().__init__()
            self.deconv = nn.ConvTranspose2d(cin, cout, k, k, 0)
            self.batch_norm = nn.BatchNorm2d(cout)

        def forward(self, x):
            return F.relu(self.batch_norm(self.deconv(x)))

    g = torch.Generator().manual_seed(cout + k)
    ref = _randomise(RefDeConv(), g)
    x = torch.randn((2, cin, H, W), generator=g).relu().to(torch.bfloat16)
    layer = vn.DeConvUpMD2d.from_reference(ref.to(cuda_device))
    assert set(layer.state_dict()) == set(ref.state_dict())
    xg = x.to(cuda_device).contiguous(memory_format=torch.channels_last)
    got = layer(xg)
    with torch.no_grad():
        want = ref.cpu().double()(x.double())
    assert got.shape == want.shape == (2, cout, H * k, W * k) and got.dtype == torch.bfloat16
    assert rel_max_err(got.cpu().double(), want) < 1e-2
    assert float((got.cpu().double() - want).norm() / want.norm()) < 5e-3
    # into the middle of a wider tensor: the neighbouring channels stay untouched
    wide = torch.full((2, H * k, W * k, cout + 128), 7.0, dtype=torch.bfloat16, device=cuda_device)
    layer.forward_nhwc(xg.permute(0, 2, 3, 1), out=wide, channel_offset=64)
    assert torch.equal(wide[..., 64:64 + cout], got.permute(0, 2, 3, 1))
    assert bool((wide[..., :64] == 7).all()) and bool((wide[..., 64 + cout:] == 7).all())


def test_rpn_block_tail_matches_the_reference_block(cuda_device):
    """A block shaped like the reference's block2 (stride-2 head + stride-1 layers, model.py:232-240), at reduced
    size: the head runs on stock PyTorch, the tail on the tcgen05 kernel, bf16 channels-last in between."""
    import voxelnet_pytorch_b200 as vn

    g = torch.Generator().manual_seed(5)
    block = nn.Sequential(RefConvMD2d(128, 128, 3, 2, 1), *[RefConvMD2d(128, 128) for _ in range(3)])
    _randomise(block, g)
    x = torch.randn((2, 128, 80, 96), generator=g).relu()
    with torch.no_grad():
        want = block.double()(x.double())
    tail = vn.RPNBlockTail(block.float().to(cuda_device))
    tail.min_tiles = 0  # (a map this small would otherwise take the stock path)
    assert [type(m).__name__ for m in tail.fast] == ["ConvMD2d", "ConvMD2d", "ConvMD2d", "ConvMD2d"]
    assert tail.on_tensor_cores(96) == [True, True, True, True]
    assert block[0].conv.weight.dtype == torch.float32  # the caller's module is left alone
    with torch.no_grad():
        got = tail(x.to(cuda_device))
    assert got.shape == (2, 128, 40, 48) and got.dtype == torch.bfloat16
    assert rel_max_err(got.cpu().double(), want) < 2e-2  # four bf16 layers in a row
    rel_l2 = float((got.cpu().double() - want).norm() / want.norm())
    assert rel_l2 < 1e-2, rel_l2


def test_front_end_bf16_output_feeds_a_conv2d_layer(cuda_device):
    """VoxelNetFrontEnd(output_format="nhwc_bf16") -> ConvMD2d with its input channels permuted once == the same layer
    with the reference's channel order on the float32 drop-in tensor rounded to bf16."""
    import voxelnet_pytorch_b200 as vn
    from oracle import model_torch as OM
    from tests.helpers import load_state, synth_frames

    sd = OM.random_state_dict(31)
    fnet, mnet = vn.FeatureLearningNet("Pedestrian").eval(), vn.MiddleConvNet().eval()
    load_state(fnet, sd, "")
    load_state(mnet, sd, "")
    fnet, mnet = fnet.to(cuda_device), mnet.to(cuda_device)
    frames = synth_frames(2, 600)
    fe32 = vn.VoxelNetFrontEnd(fnet, mnet, "Pedestrian")
    fe16 = vn.VoxelNetFrontEnd(fnet, mnet, "Pedestrian", output_format="nhwc_bf16")
    g = torch.Generator().manual_seed(6)
    ref = _randomise(RefConvMD2d(128, 128), g).to(cuda_device)
    a = vn.ConvMD2d.from_reference(ref)(fe32.forward_frames(frames).to(torch.bfloat16))
    perm = fe16.reference_channel_of_output().to(cuda_device)
    ref_p = RefConvMD2d(128, 128).to(cuda_device)
    ref_p.load_state_dict(ref.state_dict())
    ref_p.conv.weight.data = ref.conv.weight.data[:, perm].contiguous()
    b = vn.ConvMD2d.from_reference(ref_p.eval())(fe16.forward_frames(frames))
    # (the K order of the two sums differs by the channel permutation: equal up to a bf16 ulp of the output)
    assert rel_max_err(a.float().cpu(), b.float().cpu()) < 5e-3


class RefRPN(nn.Module):
    """The RPN half of the reference's MiddleConvNet for Car (model.py:212-254, forward :263-281), same attribute names."""

    def __init__(self):
        super().__init__()

        class DeConv(nn.Module):
            def __init__(self, cin, cout, k, s, p):
                super().__init__()
                self.deconv = nn.ConvTranspose2d(cin, cout, k, s, p)
                self.batch_norm = nn.BatchNorm2d(cout)

            def forward(self, x):
                return F.relu(self.batch_norm(self.deconv(x)))

        class Head(nn.Module):
            def __init__(self, cin, cout):
                super().__init__()
                self.conv = nn.Conv2d(cin, cout, 1, 1, 0)

            def forward(self, x):
                return self.conv(x)

        self.block1 = nn.Sequential(RefConvMD2d(128, 128, 3, 2, 1), *[RefConvMD2d(128, 128) for _ in range(4)])
        self.deconv1 = DeConv(128, 256, 3, 1, 1)
        self.block2 = nn.Sequential(RefConvMD2d(128, 128, 3, 2, 1), *[RefConvMD2d(128, 128) for _ in range(5)])
        self.deconv2 = DeConv(128, 256, 2, 2, 0)
        self.block3 = nn.Sequential(RefConvMD2d(128, 256, 3, 2, 1), *[RefConvMD2d(256, 256) for _ in range(5)])
        self.deconv3 = DeConv(256, 256, 4, 4, 0)
        self.prob_conv, self.reg_conv = Head(768, 2), Head(768, 14)

    def forward(self, x):
        x = self.block1(x)
        d1 = self.deconv1(x)
        x = self.block2(x)
        d2 = self.deconv2(x)
        x = self.block3(x)
        d3 = self.deconv3(x)
        x = torch.cat([d3, d2, d1], dim=1)
        return torch.sigmoid(self.prob_conv(x)), self.reg_conv(x)


def test_hybrid_rpn_matches_the_fp32_rpn(cuda_device):
    """HybridRPN (block1 / block2 tails and deconv1 on the tcgen05 kernel, the rest stock bf16) against the float32 RPN of
    the same parameters: 20 bf16 layers deep, so the bar is rel-L2 <= 2e-2 on the regression map and 2e-2 absolute on
    the probabilities; the all-stock bf16 RPN is held to the same bar as a yardstick."""
    import voxelnet_pytorch_b200 as vn

    g = torch.Generator().manual_seed(9)
    ref = _randomise(RefRPN(), g)
    x = torch.randn((2, 128, 128, 96), generator=g).relu()
    with torch.no_grad():
        p_want, r_want = ref(x)
    rpn = vn.HybridRPN(ref.to(cuda_device))
    rpn.set_min_tiles(0)  # (maps this small would otherwise take the stock path)
    rpn.heads_on_tensor_cores = True  # (the 1x1 heads as a pointwise launch too: every layer of the RPN on the kernel)
    assert all(rpn.block1.on_tensor_cores(96)) and all(rpn.block2.on_tensor_cores(48))
    assert all(rpn.block3.on_tensor_cores(24))  # 256 channels, W = 12: ragged tile column
    assert isinstance(rpn.deconv_fast[0], vn.DeConvMD2d) and isinstance(rpn.deconv_fast[1], vn.DeConvUpMD2d) \
        and isinstance(rpn.deconv_fast[2], vn.DeConvUpMD2d)
    with torch.no_grad():
        p_got, r_got = rpn(x.to(cuda_device))
        stock = ref.to(dtype=torch.bfloat16, memory_format=torch.channels_last)
        p_st, r_st = stock(x.to(cuda_device).to(torch.bfloat16).contiguous(memory_format=torch.channels_last))
        rpn.set_min_tiles(None)
        rpn.heads_on_tensor_cores = False
        p_auto, r_auto = rpn(x.to(cuda_device))  # default policy: these small maps run on the stock copies
    assert p_got.shape == (2, 2, 64, 48) and r_got.shape == (2, 14, 64, 48) and p_got.dtype == torch.float32
    for name, got in (("hybrid", (p_got, r_got)), ("hybrid, default tile policy", (p_auto, r_auto)), ("stock bf16", (p_st.float(), r_st.float()))):
        p, r = got[0].cpu(), got[1].cpu()
        assert float((p - p_want).abs().max()) < 2e-2, name
        assert float((r - r_want).norm() / r_want.norm()) < 2e-2, name


def test_middle_rpn_drop_in_forward_and_fused_inference_pair(cuda_device):
    """vn.MiddleRPN (the reference's whole MiddleConvNet under its names): forward(x (B,D,H,W,128)) = the tcgen05 3-D layers
    followed by the stock fp32 RPN half - the same bits as vn.RPN(vn.MiddleConvNet(x)) with the same parameters - and the
    inference pair `front_end()` (fused eval step, bf16 channels-last hand-over) + HybridRPN stays within the bf16 bar of
    the float32 RPN fed by the float32 front end."""
    import voxelnet_pytorch_b200 as vn
    from voxelnet_pytorch_b200 import synth

    g = torch.Generator().manual_seed(21)
    net = _randomise(vn.MiddleRPN("Pedestrian"), g).to(cuda_device).eval()
    mid, rpn32 = vn.MiddleConvNet().to(cuda_device).eval(), vn.RPN("Pedestrian").to(cuda_device).eval()
    sd = net.state_dict()
    mid.load_state_dict({k: v for k, v in sd.items() if k.startswith("middle_layer.")}, strict=True)
    rpn32.load_state_dict({k: v for k, v in sd.items() if not k.startswith("middle_layer.")}, strict=True)
    # the module forward on a small sparse grid
    dense = torch.zeros((1, 10, 32, 48, 128))
    cells = torch.randint(0, 10 * 32 * 48, (400,), generator=g)
    dense.view(-1, 128)[cells] = torch.rand((400, 128), generator=g)
    with torch.no_grad():
        p_mod, r_mod = net(dense.to(cuda_device))
        p_two, r_two = rpn32(mid(dense.to(cuda_device)))
    assert p_mod.shape == (1, 2, 32, 48) and r_mod.shape == (1, 14, 32, 48)
    assert torch.equal(p_mod, p_two) and torch.equal(r_mod, r_two)
    # the inference pair against the float32 path, from points
    fnet = vn.FeatureLearningNet("Pedestrian").to(cuda_device).eval()
    frames = [f[::3].copy() for f in synth.lidar_batch(2, seed0=31)]
    with torch.no_grad():
        y = vn.VoxelNetFrontEnd(fnet, net, "Pedestrian").forward_frames(frames).clone()  # (B,128,H,W) float32
        p_want, r_want = rpn32(y)
        fe, hy = net.front_end(fnet)
        p_got, r_got = hy(fe.forward_frames(frames))
    assert p_got.shape == p_want.shape == (2, 2, 200, 240) and r_got.shape == (2, 14, 200, 240)
    assert float((p_got - p_want).abs().max()) < 2e-2 and float((r_got - r_want).norm() / r_want.norm()) < 2e-2


def test_hybrid_rpn_matches_the_reference_rpn_golden(cuda_device):
    """HybridRPN against tests/golden/rpn_car.npz - outputs of the UNMODIFIED reference's RPN half (oracle/make_golden_rpn.py),
    parameters re-created here by the same name-keyed seeded fill.  Bars as for the fp32 comparison above (20 bf16 layers):
    regression map rel-L2 2e-2; probabilities 3e-2 absolute (the seeded He-scale weights give logits up to |8|, and at this small
    map the default tile policy = the all-stock cuDNN bf16 path was observed at 2.1e-2); every layer on the tcgen05 kernel, the
    heads too, and the default tile policy."""
    import voxelnet_pytorch_b200 as vn
    from oracle import model_torch as OM
    from tests import golden_util as G

    g = G.load("rpn_car.npz")
    seed = int(g["seed"])
    net = vn.RPN("Car").eval()
    assert OM.fill_rpn_parameters(net, seed) == str(g["params_sha256"])
    x = OM.rpn_golden_input(seed, tuple(int(v) for v in g["input_shape"])).to(cuda_device)
    want_p, want_r = torch.from_numpy(g["prob"]), torch.from_numpy(g["reg"])
    rpn = vn.HybridRPN(net.to(cuda_device))
    with torch.no_grad():
        rpn.set_min_tiles(0)
        p0, r0 = rpn(x)
        rpn.heads_on_tensor_cores = True
        p1, r1 = rpn(x)
        rpn.heads_on_tensor_cores = False
        rpn.set_min_tiles(None)
        p2, r2 = rpn(x)
    for name, (p, r) in (("all 3x3 / deconv layers on tcgen05", (p0, r0)), ("+ heads on tcgen05", (p1, r1)), ("default tile policy", (p2, r2))):
        assert p.shape == want_p.shape and r.shape == want_r.shape and p.dtype == torch.float32, name
        ep, er = float((p.cpu() - want_p).abs().max()), float((r.cpu() - want_r).norm() / want_r.norm())
        print("rpn golden: %s: prob max-abs err %.4f, reg rel-L2 %.5f" % (name, ep, er))
        assert ep < 3e-2, name
        assert er < 2e-2, name
