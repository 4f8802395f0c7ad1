"""The stride-1 3x3 layers of the RPN blocks on the tcgen05 conv kernel (second half).

Reference: `ConvMD(2, 128, 128, 3, (1, 1), (1, 1))` = Conv2d -> BatchNorm2d -> ReLU (voxelnet/model.py:111-167), the four
/ five stride-1 layers that follow the stride-2 head of `block1` and `block2` (model.py:212-242): 9 of the RPN's 17
Conv2d layers and 61 of its ~110 GFLOP per frame.  `ConvMD2d` keeps the reference's parameter names (`conv.weight`,
`conv.bias`, `batch_norm.*`), so a reference `state_dict()` loads.  Eval mode only; bf16 channels-last activations:
`forward(x)` takes and returns a `(B,C,H,W)` bf16 tensor in `torch.channels_last` memory format - what
`VoxelNetFrontEnd(output_format="nhwc_bf16")` hands over - so a block chains without a cast or layout pass.

`DeConvMD2d` is the reference's `deconv1 = DeConv2d(128, 256, 3, (1,1), (1,1))` (model.py:229; 21 GFLOP per frame): a
stride-1 transposed convolution is a convolution with the flipped kernel.  Together: 75 of the RPN's ~110 GFLOP.

The stride-2 layers, the `256`-channel block 3 (W = 44 is not a multiple of the 8-wide tile), the two strided
`ConvTranspose2d` and the two 1x1 heads stay stock PyTorch (`north_star` keeps the whole RPN there); `RPNBlockTail`
wraps a reference block so that its head runs on stock PyTorch and its stride-1 tail here.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops
from .model import _fold_bn


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

    def forward_nhwc(self, x_nhwc: torch.Tensor, out=None, workspace=None) -> torch.Tensor:
        """(B,H,W,cin) bf16 contiguous -> (B,H,W,cout) bf16."""
        if self.training:
            raise NotImplementedError("%s implements eval-mode BatchNorm (inference); train the RPN on stock PyTorch" % type(self).__name__)
        B, H, W, _ = x_nhwc.shape
        if out is None:
            out = torch.empty((B, H, W, self.cout), dtype=torch.bfloat16, device=x_nhwc.device)
        if workspace is None:
            workspace = torch.empty(ops.conv3d_workspace_bytes(B, H, W), dtype=torch.uint8, device=x_nhwc.device)
        for o, packed, shift in self.folded():
            ops.conv2d_bn_relu_fwd(x_nhwc, packed, shift, out=out, workspace=workspace, channel_offset=o)
        return out

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """(B,cin,H,W) bf16, channels_last memory -> (B,cout,H,W) bf16, channels_last memory."""
        if x.dtype != torch.bfloat16 or not x.is_contiguous(memory_format=torch.channels_last):
            x = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        return self.forward_nhwc(x.permute(0, 2, 3, 1)).permute(0, 3, 1, 2)


class ConvMD2d(_Conv3x3BnRelu):
    """model.py:111-167, 2-D case, kernel 3, stride 1, padding 1, BatchNorm + ReLU; cin in {64,128}, cout a multiple of 64."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        if cin not in (64, 128) or cout % 64 != 0 or cout < 64:
            raise ValueError("ConvMD2d: cin=%d, cout=%d unsupported by the B200 kernel (cin 64/128, cout a multiple of 64)" % (cin, cout))
        self.cin, self.cout = cin, cout
        self.conv = nn.Conv2d(cin, cout, 3, 1, 1)
        self.batch_norm = nn.BatchNorm2d(cout)

    @classmethod
    def from_reference(cls, layer) -> "ConvMD2d":
        """A reference `ConvMD(2, ...)` (or anything with `.conv` Conv2d and `.batch_norm` BatchNorm2d) -> ConvMD2d sharing
        its parameter values."""
        conv = getattr(layer, "conv", None)
        if not isinstance(conv, nn.Conv2d) or tuple(conv.kernel_size) != (3, 3) or tuple(conv.stride) != (1, 1) or tuple(conv.padding) != (1, 1):
            raise ValueError("ConvMD2d covers 3x3 / stride 1 / padding 1 Conv2d layers only")
        m = cls(conv.in_channels, conv.out_channels)
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
        if cin not in (64, 128) or cout % 64 != 0 or cout < 64:
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


class RPNBlockTail(nn.Module):
    """A reference RPN block (`nn.Sequential` of ConvMD(2, ...), model.py:212-251) with its stride-1 3x3 layers on the
    tcgen05 kernel; layers the kernel does not cover (the stride-2 head) run as they are, in bf16 channels-last."""

    def __init__(self, block: nn.Sequential):
        super().__init__()
        layers = []
        for layer in block:
            try:
                layers.append(ConvMD2d.from_reference(layer))
            except ValueError:
                layers.append(layer.to(dtype=torch.bfloat16, memory_format=torch.channels_last))
        self.layers = nn.ModuleList(layers)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        ws = None
        for layer in self.layers:
            if isinstance(layer, ConvMD2d):
                B, _, H, W = x.shape
                if W % 8 != 0:
                    raise ValueError("ConvMD2d needs W % 8 == 0; got W=%d" % W)
                if ws is None or ws[0] != (B, H, W):
                    ws = ((B, H, W), torch.empty(ops.conv3d_workspace_bytes(B, H, W), dtype=torch.uint8, device=x.device))
                x = layer.forward_nhwc(x.permute(0, 2, 3, 1), workspace=ws[1]).permute(0, 3, 1, 2)
            else:
                x = layer(x)
        return x
