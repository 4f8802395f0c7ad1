"""The 3x3 layers of the RPN blocks (stride 1 and 2) and deconv1 on the tcgen05 conv kernel (second half).

Reference: `ConvMD(2, 128, 128, 3, (1, 1), (1, 1))` = Conv2d -> BatchNorm2d -> ReLU (voxelnet/model.py:111-167), the four
/ five stride-1 layers that follow the stride-2 head of `block1` and `block2` (model.py:212-242): 9 of the RPN's 17
Conv2d layers and 61 of its ~110 GFLOP per frame.  `ConvMD2d` keeps the reference's parameter names (`conv.weight`,
`conv.bias`, `batch_norm.*`), so a reference `state_dict()` loads.  Eval mode only; bf16 channels-last activations:
`forward(x)` takes and returns a `(B,C,H,W)` bf16 tensor in `torch.channels_last` memory format - what
`VoxelNetFrontEnd(output_format="nhwc_bf16")` hands over - so a block chains without a cast or layout pass.

`DeConvMD2d` is the reference's `deconv1 = DeConv2d(128, 256, 3, (1,1), (1,1))` (model.py:229; 21 GFLOP per frame): a
stride-1 transposed convolution is a convolution with the flipped kernel.  The five 256 -> 256 layers of block3 run as
two launches of 128 output channels over four 64-channel input chunks (W = 44: the last tile column is ragged).
Together 15 of the RPN's 22 layers, 88 of its ~110 GFLOP.

The stride-2 heads of the blocks run on the kernel's stride-2 instantiation (odd / even input rows as two boxes).
`DeConvUpMD2d` is deconv2 / deconv3 (kernel = stride 2 / 4): k*k 1x1 launches that write each output pixel once.
The two 1x1 heads (no BatchNorm, no activation) can run as one pointwise launch padded to 64 output channels
(`HybridRPN.heads_on_tensor_cores`; the fused stock 1x1 convolution is 2 % faster and is the default) (`north_star` keeps the whole RPN there); `RPNBlockTail`
wraps a reference block so that its head runs on stock PyTorch and its stride-1 tail here, `HybridRPN` the whole RPN half
of the reference's `MiddleConvNet`.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops
from .config import grid_config
from .model import _fold_bn


class _PlanCache:
    """Workspaces of the dense 2-D launches by (B, H, W, output channels of a launch): the work plan of such a launch depends
    on nothing else, so the second and later launches of that shape reuse it (`planned`)."""

    def __init__(self):
        self._ws = {}

    def get(self, B, H, W, n, device, kind: str = "3x3"):
        key = (B, H, W, n, str(device), kind)
        hit = key in self._ws
        if not hit:
            self._ws[key] = torch.empty(ops.conv2d_workspace_bytes(B, H, W), dtype=torch.uint8, device=device)
        return self._ws[key], hit


def _enough_tiles(B: int, H: int, W: int, device, min_tiles=None) -> bool:
    """The persistent conv kernel wants a few tiles per SM; tiny feature maps (block3 at batch 16: 384 tiles on 148 SMs)
    are faster on the stock path.  min_tiles: threshold (default 4 per SM; 0 forces the tcgen05 path)."""
    if min_tiles is None:
        min_tiles = 4 * (torch.cuda.get_device_properties(device).multi_processor_count if torch.cuda.is_available() else 148)
    return B * ((H + 15) // 16) * ((W + 7) // 8) >= min_tiles


class _Conv3x3BnRelu(nn.Module):
    """Shared eval-mode path: a 3x3 / stride 1 / padding 1 convolution in torch Conv2d weight form + BatchNorm2d + ReLU,
    split into launches of <= 192 output channels that write their slice of one channels-last output tensor."""

    cin: int
    cout: int
    batch_norm: nn.BatchNorm2d

    def _conv_weight_bias(self):
        raise NotImplementedError

    def _tracked(self):
        raise NotImplementedError

    def folded(self):
        """[(first output channel, packed bf16 weights with the BatchNorm scale folded in, shift (n,) f32), ...]."""
        ts = self._tracked() + [self.batch_norm.weight, self.batch_norm.bias, self.batch_norm.running_mean, self.batch_norm.running_var]
        key = tuple((t.data_ptr(), t._version) for t in ts)
        if getattr(self, "_packed", None) is None or self._key != key:
            w, bias = self._conv_weight_bias()
            scale, bn_shift = _fold_bn(self.batch_norm)
            shift = (bias.detach().double() * scale.double() + bn_shift.double()).float().contiguous()
            parts, o = [], 0
            while o < self.cout:
                n = min(128 if self.cout - o != 192 else 192, self.cout - o)
                parts.append((o, ops.conv2d_pack_weights(w[o:o + n].contiguous(), scale[o:o + n].contiguous()), shift[o:o + n].contiguous()))
                o += n
            self._packed, self._key = parts, key
        return self._packed

    def forward_nhwc(self, x_nhwc: torch.Tensor, out=None, plans: "_PlanCache" = None, channel_offset: int = 0) -> torch.Tensor:
        """(B,H,W,cin) bf16 contiguous -> (B,H,W,cout) bf16 (or channels channel_offset .. + cout of a wider `out`)."""
        if self.training:
            raise NotImplementedError("%s implements eval-mode BatchNorm (inference); train the RPN on stock PyTorch" % type(self).__name__)
        B, H, W, _ = x_nhwc.shape
        stride = int(getattr(self, "stride", 1))
        if stride == 2:
            H, W = (H - 1) // 2 + 1, (W - 1) // 2 + 1
        if out is None:
            out = torch.empty((B, H, W, self.cout), dtype=torch.bfloat16, device=x_nhwc.device)
        if plans is None:
            plans = self.__dict__.setdefault("_plans", _PlanCache())
        for o, packed, shift in self.folded():
            ws, planned = plans.get(B, H, W, int(shift.numel()), x_nhwc.device)
            ops.conv2d_bn_relu_fwd(x_nhwc, packed, shift, out=out, workspace=ws, channel_offset=channel_offset + o, planned=planned,
                                   stride=stride)
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """(B,cin,H,W) bf16, channels_last memory -> (B,cout,H,W) bf16, channels_last memory."""
        if x.dtype != torch.bfloat16 or not x.is_contiguous(memory_format=torch.channels_last):
            x = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        return self.forward_nhwc(x.permute(0, 2, 3, 1)).permute(0, 3, 1, 2)


class ConvMD2d(_Conv3x3BnRelu):
    """model.py:111-167, 2-D case, kernel 3, stride 1 or 2, padding 1, BatchNorm + ReLU; cin in {64,128,256}, cout a multiple of 64."""

    def __init__(self, cin: int, cout: int, stride: int = 1):
        super().__init__()
        if cin not in (64, 128, 256) or cout % 64 != 0 or cout < 64 or stride not in (1, 2):
            raise ValueError("ConvMD2d: cin=%d, cout=%d, stride=%d unsupported by the B200 kernel (cin 64/128/256, cout a multiple "
                             "of 64, stride 1/2)" % (cin, cout, stride))
        self.cin, self.cout, self.stride = cin, cout, stride
        self.conv = nn.Conv2d(cin, cout, 3, stride, 1)
        self.batch_norm = nn.BatchNorm2d(cout)

    @classmethod
    def from_reference(cls, layer) -> "ConvMD2d":
        """A reference `ConvMD(2, ...)` (or anything with `.conv` Conv2d and `.batch_norm` BatchNorm2d) -> ConvMD2d sharing
        its parameter values."""
        conv = getattr(layer, "conv", None)
        if not isinstance(conv, nn.Conv2d) or tuple(conv.kernel_size) != (3, 3) or tuple(conv.stride) not in ((1, 1), (2, 2)) \
                or tuple(conv.padding) != (1, 1):
            raise ValueError("ConvMD2d covers 3x3 / stride 1 or 2 / padding 1 Conv2d layers only")
        m = cls(conv.in_channels, conv.out_channels, int(conv.stride[0]))
        m.load_state_dict(layer.state_dict())
        return m.to(conv.weight.device).eval()

    def _conv_weight_bias(self):
        return self.conv.weight.detach().float(), self.conv.bias

    def _tracked(self):
        return [self.conv.weight, self.conv.bias]


class DeConvMD2d(_Conv3x3BnRelu):
    """model.py:170-199 for kernel 3, stride 1, padding 1 (`deconv1 = DeConv2d(128, 256, 3, (1,1), (1,1))`, model.py:229):
    ConvTranspose2d -> BatchNorm2d -> ReLU.  A stride-1 transposed convolution is the convolution with the kernel flipped
    and its channel axes swapped, so it runs on the same kernel.  Parameter names as in the reference (`deconv.*`,
    `batch_norm.*`)."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        if cin not in (64, 128, 256) or cout % 64 != 0 or cout < 64:
            raise ValueError("DeConvMD2d: cin=%d, cout=%d unsupported by the B200 kernel" % (cin, cout))
        self.cin, self.cout = cin, cout
        self.deconv = nn.ConvTranspose2d(cin, cout, 3, 1, 1)
        self.batch_norm = nn.BatchNorm2d(cout)

    @classmethod
    def from_reference(cls, layer) -> "DeConvMD2d":
        d = getattr(layer, "deconv", None)
        if not isinstance(d, nn.ConvTranspose2d) or tuple(d.kernel_size) != (3, 3) or tuple(d.stride) != (1, 1) or tuple(d.padding) != (1, 1) \
                or tuple(d.output_padding) != (0, 0):
            raise ValueError("DeConvMD2d covers 3x3 / stride 1 / padding 1 ConvTranspose2d layers only")
        m = cls(d.in_channels, d.out_channels)
        m.load_state_dict(layer.state_dict())
        return m.to(d.weight.device).eval()

    def _conv_weight_bias(self):
        return self.deconv.weight.detach().float().permute(1, 0, 2, 3).flip(2, 3).contiguous(), self.deconv.bias

    def _tracked(self):
        return [self.deconv.weight, self.deconv.bias]


class DeConvUpMD2d(nn.Module):
    """model.py:170-199 for kernel = stride = k, padding 0 (`deconv2 = DeConv2d(128, 256, 2, (2,2), (0,0))`, `deconv3 =
    DeConv2d(256, 256, 4, (4,4), (0,0))`, model.py:242,253): every output pixel (k y + i, k x + j) is a 1x1 convolution of
    input pixel (y, x) with W[:, :, i, j] - k*k launches of `vn_conv2d_pointwise_fwd` per block of 128 output channels, each
    writing its pixels once, BatchNorm + ReLU in the epilogue.  Parameter names as in the reference."""

    def __init__(self, cin: int, cout: int, k: int):
        super().__init__()
        if cin not in (64, 128, 256) or cout % 64 != 0 or cout < 64 or not 1 <= k <= 8:
            raise ValueError("DeConvUpMD2d: cin=%d, cout=%d, k=%d unsupported by the B200 kernel" % (cin, cout, k))
        self.cin, self.cout, self.k = cin, cout, k
        self.deconv = nn.ConvTranspose2d(cin, cout, k, k, 0)
        self.batch_norm = nn.BatchNorm2d(cout)

    @classmethod
    def from_reference(cls, layer) -> "DeConvUpMD2d":
        d = getattr(layer, "deconv", None)
        if not isinstance(d, nn.ConvTranspose2d) or d.kernel_size[0] != d.kernel_size[1] or tuple(d.stride) != tuple(d.kernel_size) \
                or tuple(d.padding) != (0, 0) or tuple(d.output_padding) != (0, 0):
            raise ValueError("DeConvUpMD2d covers ConvTranspose2d with kernel = stride and no padding only")
        m = cls(d.in_channels, d.out_channels, int(d.kernel_size[0]))
        m.load_state_dict(layer.state_dict())
        return m.to(d.weight.device).eval()

    def folded(self):
        """[(oy, ox, first output channel, packed 1x1 weights with the BatchNorm scale folded in, shift (n,) f32), ...]."""
        ts = [self.deconv.weight, self.deconv.bias, self.batch_norm.weight, self.batch_norm.bias, self.batch_norm.running_mean,
              self.batch_norm.running_var]
        key = tuple((t.data_ptr(), t._version) for t in ts)
        if getattr(self, "_packed", None) is None or self._key != key:
            scale, bn_shift = _fold_bn(self.batch_norm)
            shift = (self.deconv.bias.detach().double() * scale.double() + bn_shift.double()).float().contiguous()
            w = self.deconv.weight.detach().float()  # (cin, cout, k, k)
            parts = []
            for oy in range(self.k):
                for ox in range(self.k):
                    w_ij = w[:, :, oy, ox].t().contiguous()[:, :, None, None]  # (cout, cin, 1, 1)
                    o = 0
                    while o < self.cout:
                        n = min(128 if self.cout - o != 192 else 192, self.cout - o)
                        parts.append((oy, ox, o, ops.conv2d_pack_weights(w_ij[o:o + n].contiguous(), scale[o:o + n].contiguous()),
                                      shift[o:o + n].contiguous()))
                        o += n
            self._packed, self._key = parts, key
        return self._packed

    def forward_nhwc(self, x_nhwc: torch.Tensor, out=None, plans: "_PlanCache" = None, channel_offset: int = 0) -> torch.Tensor:
        """(B,H,W,cin) bf16 contiguous -> (B,kH,kW,cout) bf16 (or channels channel_offset .. + cout of a wider `out`)."""
        if self.training:
            raise NotImplementedError("DeConvUpMD2d implements eval-mode BatchNorm (inference)")
        B, H, W, _ = x_nhwc.shape
        if out is None:
            out = torch.empty((B, H * self.k, W * self.k, self.cout), dtype=torch.bfloat16, device=x_nhwc.device)
        if plans is None:
            plans = self.__dict__.setdefault("_plans", _PlanCache())
        for oy, ox, o, packed, shift in self.folded():
            ws, planned = plans.get(B, H, W, int(shift.numel()), x_nhwc.device, "1x1")
            ops.conv2d_pointwise_fwd(x_nhwc, packed, shift, out, channel_offset=channel_offset + o, up=self.k, oy=oy, ox=ox,
                                     workspace=ws, planned=planned)
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.forward_nhwc(_as_bf16_nhwc(x).permute(0, 2, 3, 1)).permute(0, 3, 1, 2)


class _FoldedStock(nn.Module):
    """A stock layer for inference: Conv2d / ConvTranspose2d with the eval-mode BatchNorm folded into its weights and bias
    (one cuDNN call + an in-place ReLU instead of three passes over the activation), bf16 channels-last."""

    def __init__(self, op: nn.Module, bn, relu: bool):
        super().__init__()
        import copy

        op = copy.deepcopy(op).float()
        if bn is not None:
            scale, shift = _fold_bn(bn)
            scale, shift = scale.to(op.weight.device), shift.to(op.weight.device)
            shape = (1, -1, 1, 1) if isinstance(op, nn.ConvTranspose2d) else (-1, 1, 1, 1)
            op.weight.data = op.weight.data * scale.view(shape)
            op.bias.data = op.bias.data * scale + shift
        self.op = op.to(dtype=torch.bfloat16, memory_format=torch.channels_last).eval()
        self.relu = relu
        self.stride = tuple(op.stride)

    @property
    def conv(self):  # (the attribute RPNBlockTail / permute_input_channels look at)
        return self.op

    def forward(self, x):
        y = self.op(x)
        return torch.relu_(y) if self.relu else y


def _stock_bf16(layer: nn.Module) -> nn.Module:
    """A private bf16 channels-last inference copy of a reference ConvMD(2, ...) / DeConv2d (model.py:111-199); the
    caller's module is left as it is."""
    op = getattr(layer, "conv", None) or getattr(layer, "deconv", None)
    if op is None:
        import copy

        return copy.deepcopy(layer).to(dtype=torch.bfloat16, memory_format=torch.channels_last).eval()
    bn = getattr(layer, "batch_norm", None) if getattr(layer, "bn", True) else None
    relu = bool(getattr(layer, "activation", True))
    return _FoldedStock(op, bn, relu)


def _as_bf16_nhwc(x: torch.Tensor) -> torch.Tensor:
    if x.dtype != torch.bfloat16 or not x.is_contiguous(memory_format=torch.channels_last):
        x = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    return x


class RPNBlockTail(nn.Module):
    """A reference RPN block (`nn.Sequential` of ConvMD(2, ...), model.py:212-251) with its stride-1 3x3 layers on the
    tcgen05 kernel; layers the kernel does not cover (the stride-2 head) run on a private bf16 channels-last copy of the
    stock layer."""

    def __init__(self, block: nn.Sequential):
        super().__init__()
        fast, stock = [], []
        for layer in block:
            try:
                fast.append(ConvMD2d.from_reference(layer))
            except ValueError:
                fast.append(None)
            stock.append(_stock_bf16(layer))
        self.stock = nn.ModuleList(stock)
        self.fast = nn.ModuleList([m if m is not None else nn.Identity() for m in fast])
        self._plans = _PlanCache()
        self.min_tiles = None  # see _enough_tiles

    def on_tensor_cores(self, width: int) -> list:
        """Which layers run on the tcgen05 kernel for an input of this width (after the block's own strides)."""
        out, w = [], width
        for f, s in zip(self.fast, self.stock):
            w = w // int(s.conv.stride[1])
            out.append(isinstance(f, ConvMD2d))
        return out

    def forward(self, x: torch.Tensor, plans: _PlanCache = None) -> torch.Tensor:
        x = _as_bf16_nhwc(x)
        plans = self._plans if plans is None else plans
        min_tiles = self.min_tiles
        for f, s in zip(self.fast, self.stock):
            B, _, H, W = x.shape
            st = int(s.conv.stride[0])
            if isinstance(f, ConvMD2d) and _enough_tiles(B, (H - 1) // st + 1, (W - 1) // st + 1, x.device, min_tiles):
                x = f.forward_nhwc(x.permute(0, 2, 3, 1), plans=plans).permute(0, 3, 1, 2)
            else:
                x = s(x)
        return x


def pc_in_ok(cat: torch.Tensor) -> bool:
    """The 1x1 heads run on the tcgen05 kernel when the concatenated tensor's channel count is a multiple of 64 (<= 1024)."""
    c = int(cat.shape[3])
    return c % 64 == 0 and 64 <= c <= 1024


class HybridRPN(nn.Module):
    """The RPN half of the reference's `MiddleConvNet` (model.py:212-254, forward :263-281) in bf16 channels-last, with the
    layers the tcgen05 kernel covers running on it - the stride-1 3x3 layers of block1 / block2 and deconv1, 75 of the
    ~110 GFLOP per Car frame - and everything else (stride-2 heads, block3, deconv2 / deconv3, the 1x1 heads, the
    concatenation, the sigmoid) on private bf16 copies of the stock PyTorch layers.

        rpn = HybridRPN(reference_model.middle_rpn)                 # eval mode; parameters are copied
        fe = VoxelNetFrontEnd(fnet, mnet, 'Car', output_format="nhwc_bf16")
        rpn.permute_input_channels(fe.reference_channel_of_output())
        prob, reg = rpn(fe.forward_frames(frames))                  # (B,2,H/2,W/2), (B,14,H/2,W/2) float32

    `middle_rpn` is anything with the reference's attribute names: block1..3 (Sequential of ConvMD(2,..)), deconv1..3
    (DeConv2d), prob_conv, reg_conv."""

    def __init__(self, middle_rpn: nn.Module):
        super().__init__()
        self.block1, self.block2, self.block3 = (RPNBlockTail(getattr(middle_rpn, n)) for n in ("block1", "block2", "block3"))
        self.deconv_stock = nn.ModuleList([_stock_bf16(getattr(middle_rpn, n)) for n in ("deconv1", "deconv2", "deconv3")])
        fast = []
        for n in ("deconv1", "deconv2", "deconv3"):
            m = nn.Identity()
            for cls in (DeConvMD2d, DeConvUpMD2d):
                try:
                    m = cls.from_reference(getattr(middle_rpn, n))
                    break
                except ValueError:
                    pass
            fast.append(m)
        self.deconv_fast = nn.ModuleList(fast)
        # the two 1x1 heads read the same 768-channel tensor: one convolution with their weights stacked (2 + 14 outputs)
        import copy

        pc, rc = middle_rpn.prob_conv.conv, middle_rpn.reg_conv.conv
        head = nn.Conv2d(pc.in_channels, pc.out_channels + rc.out_channels, 1, 1, 0)
        head.weight.data = torch.cat([pc.weight.data, rc.weight.data], 0).clone()
        head.bias.data = torch.cat([pc.bias.data, rc.bias.data], 0).clone()
        self.heads = head.to(device=pc.weight.device, dtype=torch.bfloat16, memory_format=torch.channels_last).eval()
        self.n_prob = pc.out_channels
        # ... and the same 1x1 convolution on the tcgen05 kernel: 16 outputs padded to one group of 64, no activation
        n_out = pc.out_channels + rc.out_channels
        w64 = torch.zeros((64, pc.in_channels, 1, 1), dtype=torch.float32, device=pc.weight.device)
        b64 = torch.zeros(64, dtype=torch.float32, device=pc.weight.device)
        w64[:n_out] = torch.cat([pc.weight.data, rc.weight.data], 0).float()
        b64[:n_out] = torch.cat([pc.bias.data, rc.bias.data], 0).float()
        self._head_w64, self._head_b64, self._head_packed = w64, b64, None
        # measured at Car batch 16: 3.07 ms with the heads on the kernel (N = 64 MMAs over 12 channel chunks) vs 3.00 ms with the
        # fused stock 1x1 convolution - off by default, parity-tested
        self.heads_on_tensor_cores = False
        self.n_out = n_out
        # the strided deconvolutions write every output once; even a map with few tiles beats the stock path's three passes
        self.deconv_min_tiles = 0

    def set_min_tiles(self, min_tiles) -> None:
        """Feature maps with fewer 16x8 tiles than this run on the stock path (None: 4 per SM; 0: always tcgen05)."""
        for b in (self.block1, self.block2, self.block3):
            b.min_tiles = min_tiles
        self.deconv_min_tiles = 0 if min_tiles is None else min_tiles

    def permute_input_channels(self, reference_channel_of_input: torch.Tensor) -> None:
        """The input's channel k holds the reference's channel reference_channel_of_input[k] (VoxelNetFrontEnd
        .reference_channel_of_output()): permute the input channels of block1's first layer once, so that the block sees
        the same convolution."""
        perm = reference_channel_of_input.to(self.block1.stock[0].conv.weight.device)
        for m in (self.block1.stock[0], self.block1.fast[0]):
            if hasattr(m, "conv"):
                m.conv.weight.data = m.conv.weight.data[:, perm].contiguous()

    def _deconv_up(self, i: int, x: torch.Tensor, cat: torch.Tensor, channel_offset: int, plans) -> None:
        """deconv2 / deconv3 into their channel slice of the concatenated tensor: k*k pointwise launches on the tcgen05 kernel
        when the map has enough tiles, else the stock copy + a slice copy."""
        f = self.deconv_fast[i]
        B, _, H, W = x.shape
        if isinstance(f, DeConvUpMD2d) and _enough_tiles(B, H, W, x.device, self.deconv_min_tiles):
            f.forward_nhwc(x.permute(0, 2, 3, 1), out=cat, plans=plans, channel_offset=channel_offset)
        else:
            y = self.deconv_stock[i](x)
            cat[..., channel_offset:channel_offset + y.shape[1]] = y.permute(0, 2, 3, 1)

    def forward(self, x: torch.Tensor):
        plans = self.__dict__.setdefault("_plans", _PlanCache())
        x = _as_bf16_nhwc(x)
        x = self.block1(x, plans)
        B, _, H, W = x.shape
        # model.py:272-274: cat([deconv3, deconv2, deconv1], dim=1) - one channels-last buffer that every branch writes its slice of
        c3, c2, c1 = (int(m.op.out_channels) for m in self.deconv_stock[::-1])
        cat = torch.empty((B, H, W, c3 + c2 + c1), dtype=torch.bfloat16, device=x.device)
        f1 = self.deconv_fast[0]
        if isinstance(f1, DeConvMD2d) and _enough_tiles(B, H, W, x.device, self.block1.min_tiles):
            f1.forward_nhwc(x.permute(0, 2, 3, 1), out=cat, plans=plans, channel_offset=c3 + c2)
        else:
            cat[..., c3 + c2:] = self.deconv_stock[0](x).permute(0, 2, 3, 1)
        x = self.block2(x, plans)
        self._deconv_up(1, x, cat, c3, plans)
        x = self.block3(x, plans)
        self._deconv_up(2, x, cat, 0, plans)
        if self.heads_on_tensor_cores and pc_in_ok(cat) and cat.is_cuda:
            if self._head_packed is None:
                self._head_packed = ops.conv2d_pack_weights(self._head_w64)
            ws, planned = plans.get(B, H, W, 64, cat.device, "1x1")
            y64 = torch.empty((B, H, W, 64), dtype=torch.bfloat16, device=cat.device)
            ops.conv2d_pointwise_fwd(cat, self._head_packed, self._head_b64, y64, workspace=ws, planned=planned, relu=False)
            y = y64[..., : self.n_out].permute(0, 3, 1, 2).float()
            return torch.sigmoid(y[:, : self.n_prob]).contiguous(), y[:, self.n_prob:].contiguous()
        x = cat.permute(0, 3, 1, 2)
        y = self.heads(x).float()
        return torch.sigmoid(y[:, : self.n_prob]).contiguous(), y[:, self.n_prob:].contiguous()


class _StockConvMD2d(nn.Module):
    """Parameter holder with the reference's ConvMD(2, ...) attribute names (`conv`, `batch_norm`, `bn`, `activation`;
    model.py:111-167) on stock PyTorch layers - the fp32 form `RPN` trains and checkpoints in."""

    def __init__(self, cin: int, cout: int, k: int, stride: int, pad: int, bn: bool = True, activation: bool = True):
        super().__init__()
        self.cin, self.cout, self.bn, self.activation = cin, cout, bn, activation
        self.conv = nn.Conv2d(cin, cout, k, (stride, stride), (pad, pad))
        if bn:
            self.batch_norm = nn.BatchNorm2d(cout)

    def forward(self, x):
        x = self.conv(x)
        if self.bn:
            x = self.batch_norm(x)
        return torch.relu(x) if self.activation else x


class _StockDeConv2d(nn.Module):
    """The reference's DeConv2d attribute names (`deconv`, `batch_norm`, `bn`; model.py:170-199) on stock layers."""

    def __init__(self, cin: int, cout: int, k: int, stride: int, pad: int):
        super().__init__()
        self.cin, self.cout, self.bn = cin, cout, True
        self.deconv = nn.ConvTranspose2d(cin, cout, k, (stride, stride), (pad, pad))
        self.batch_norm = nn.BatchNorm2d(cout)

    def forward(self, x):
        return torch.relu(self.batch_norm(self.deconv(x)))


class RPN(nn.Module):
    """The region-proposal half of the reference's `MiddleConvNet` (model.py:212-254; forward :263-281) under the
    reference's parameter names (block1..3, deconv1..3, prob_conv, reg_conv), so that the RPN entries of a reference
    checkpoint load with a strict `load_state_dict` once the `middle_layer.*` keys are dropped.  This class is
    the fp32 stock form (training, checkpoints, the parity reference); `HybridRPN(RPN(...))` is the inference form on the
    tcgen05 kernel.  Car: block1 starts with a stride-2 layer; Pedestrian / Cyclist: stride 1 (model.py:220-236)."""

    def __init__(self, cls_name: str = "Car"):
        super().__init__()
        s1 = 2 if cls_name == "Car" else 1

        def block(cin, cout, n, stride):
            return nn.Sequential(_StockConvMD2d(cin, cout, 3, stride, 1), *[_StockConvMD2d(cout, cout, 3, 1, 1) for _ in range(n - 1)])

        self.block1 = block(128, 128, 5, s1)
        self.deconv1 = _StockDeConv2d(128, 256, 3, 1, 1)
        self.block2 = block(128, 128, 6, 2)
        self.deconv2 = _StockDeConv2d(128, 256, 2, 2, 0)
        self.block3 = block(128, 256, 6, 2)
        self.deconv3 = _StockDeConv2d(256, 256, 4, 4, 0)
        self.prob_conv = _StockConvMD2d(768, 2, 1, 1, 0, bn=False, activation=False)
        self.reg_conv = _StockConvMD2d(768, 14, 1, 1, 0, bn=False, activation=False)
        g = grid_config(cls_name)
        ratio = 2 if cls_name == "Car" else 1  # config.py:51-53 / :74-76 (FEATURE_RATIO)
        self.output_shape = [g.height // ratio, g.width // ratio]

    def forward(self, x: torch.Tensor):
        """x: (B, 128, H, W) float32 - the front end's map.  Returns (sigmoid(prob) (B,2,h,w), reg (B,14,h,w))."""
        x = self.block1(x)
        d1 = self.deconv1(x)
        x = self.block2(x)
        d2 = self.deconv2(x)
        x = self.block3(x)
        d3 = self.deconv3(x)
        x = torch.cat([d3, d2, d1], dim=1)
        return torch.sigmoid(self.prob_conv(x)), self.reg_conv(x)


class MiddleRPN(nn.Module):
    """Drop-in for the reference's WHOLE `MiddleConvNet` (model.py:202-281): the three 3-D layers (`middle_layer`, on the
    tcgen05 kernel like `MiddleConvNet` here) and the RPN half (`block1..3`, `deconv1..3`, `prob_conv`, `reg_conv`) under
    the reference's attribute and parameter names, so `RPN3D.middle_rpn = MiddleRPN(cls_name)` keeps a reference checkpoint
    loading with a strict `load_state_dict`.  forward(x (B,D,H,W,128)) -> (sigmoid(prob) (B,2,h,w), reg (B,14,h,w)) like
    model.py:257-281; the RPN half of this forward is the stock fp32 path (training, parity).

    Inference: `front_end(feature_net, cls_name)` + `hybrid()` give the fused eval step and the bf16 tcgen05 RPN that
    belong together (channel permutation of block1's first layer applied)."""

    _RPN_PARTS = ("block1", "deconv1", "block2", "deconv2", "block3", "deconv3", "prob_conv", "reg_conv")

    def __init__(self, cls_name: str = "Car"):
        super().__init__()
        from .model import MiddleConvNet

        self.cls_name = cls_name
        self.middle_layer = MiddleConvNet().middle_layer
        rpn = RPN(cls_name)
        for name in self._RPN_PARTS:  # registration order = the reference's (state_dict order matches too)
            setattr(self, name, getattr(rpn, name))
        self.output_shape = rpn.output_shape

    def forward(self, x: torch.Tensor):
        from .model import MiddleConvNet

        return RPN.forward(self, MiddleConvNet.forward(self, x))  # (both only read the sub-modules registered above)

    def hybrid(self) -> "HybridRPN":
        """The RPN half as `HybridRPN` (a snapshot of the current parameters; eval mode)."""
        return HybridRPN(self)

    def front_end(self, feature_net: nn.Module, **kw):
        """(`VoxelNetFrontEnd` over feature_net + this module's 3-D layers with bf16 channels-last output, `HybridRPN` whose
        first layer takes that tensor's channel order): `prob, reg = rpn(fe.forward_frames(frames))`."""
        from .pipeline import VoxelNetFrontEnd

        fe = VoxelNetFrontEnd(feature_net, self, self.cls_name, output_format="nhwc_bf16", **kw)
        rpn = self.hybrid()
        rpn.permute_input_channels(fe.reference_channel_of_output())
        return fe, rpn
