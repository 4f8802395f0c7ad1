"""The stride-1 3x3 layers of the RPN blocks on the tcgen05 conv kernel (second half).

Reference: `ConvMD(2, 128, 128, 3, (1, 1), (1, 1))` = Conv2d -> BatchNorm2d -> ReLU (voxelnet/model.py:111-167), the four
/ five stride-1 layers that follow the stride-2 head of `block1` and `block2` (model.py:212-242): 9 of the RPN's 17
Conv2d layers and 61 of its ~110 GFLOP per frame.  `ConvMD2d` keeps the reference's parameter names (`conv.weight`,
`conv.bias`, `batch_norm.*`), so a reference `state_dict()` loads.  Eval mode only; bf16 channels-last activations:
`forward(x)` takes and returns a `(B,C,H,W)` bf16 tensor in `torch.channels_last` memory format - what
`VoxelNetFrontEnd(output_format="nhwc_bf16")` hands over - so a block chains without a cast or layout pass.

The stride-2 layers, the `256`-channel block 3 (W = 44 is not a multiple of the 8-wide tile), the three
`ConvTranspose2d` and the two 1x1 heads stay stock PyTorch (`north_star` keeps the whole RPN there); `RPNBlockTail`
wraps a reference block so that its head runs on stock PyTorch and its stride-1 tail here.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops
from .model import _fold_bn


class ConvMD2d(nn.Module):
    """model.py:111-167, 2-D case, kernel 3, stride 1, padding 1, BatchNorm + ReLU, cin in {64,128}, cout in {64,128,192}."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        if cin not in (64, 128) or cout not in (64, 128, 192):
            raise ValueError("ConvMD2d: cin=%d, cout=%d unsupported by the B200 kernel (cin 64/128, cout 64/128/192)" % (cin, cout))
        self.cin, self.cout = cin, cout
        self.conv = nn.Conv2d(cin, cout, 3, 1, 1)
        self.batch_norm = nn.BatchNorm2d(cout)
        self._packed, self._key = None, None

    @classmethod
    def from_reference(cls, layer) -> "ConvMD2d":
        """A reference `ConvMD(2, ...)` (or anything with `.conv` Conv2d and `.batch_norm` BatchNorm2d) -> ConvMD2d sharing
        its parameter values."""
        conv = layer.conv
        if tuple(conv.kernel_size) != (3, 3) or tuple(conv.stride) != (1, 1) or tuple(conv.padding) != (1, 1):
            raise ValueError("ConvMD2d covers 3x3 / stride 1 / padding 1 layers only")
        m = cls(conv.in_channels, conv.out_channels)
        m.load_state_dict({k: v for k, v in layer.state_dict().items()})
        return m.to(conv.weight.device).eval()

    def folded(self):
        """(packed bf16 weights with the BatchNorm scale folded in, shift (cout,) f32)."""
        ts = [self.conv.weight, self.conv.bias, self.batch_norm.weight, self.batch_norm.bias, self.batch_norm.running_mean,
              self.batch_norm.running_var]
        key = tuple((t.data_ptr(), t._version) for t in ts)
        if self._packed is None or self._key != key:
            scale, bn_shift = _fold_bn(self.batch_norm)
            shift = (self.conv.bias.detach().double() * scale.double() + bn_shift.double()).float().contiguous()
            self._packed = (ops.conv2d_pack_weights(self.conv.weight, scale), shift)
            self._key = key
        return self._packed

    def forward_nhwc(self, x_nhwc: torch.Tensor, out=None, workspace=None) -> torch.Tensor:
        """(B,H,W,cin) bf16 contiguous -> (B,H,W,cout) bf16."""
        if self.training:
            raise NotImplementedError("ConvMD2d implements eval-mode BatchNorm (inference); train the RPN on stock PyTorch")
        packed, shift = self.folded()
        return ops.conv2d_bn_relu_fwd(x_nhwc, packed, shift, out=out, workspace=workspace)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """(B,cin,H,W) bf16, channels_last memory -> (B,cout,H,W) bf16, channels_last memory."""
        if x.dtype != torch.bfloat16 or not x.is_contiguous(memory_format=torch.channels_last):
            x = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        return self.forward_nhwc(x.permute(0, 2, 3, 1)).permute(0, 3, 1, 2)


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
