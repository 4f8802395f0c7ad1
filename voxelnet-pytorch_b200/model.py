"""The reference's front-end modules, re-implemented on libvoxelnet_b200 kernels.

Mirrors (same class names, constructor arguments, forward signatures, parameter names):

* `VFELayer`            voxelnet/model.py:60-82
* `FeatureLearningNet`  voxelnet/model.py:85-108
* `ConvMD` (3-D only)   voxelnet/model.py:111-167
* `MiddleConvNet`       voxelnet/model.py:202-262, up to and including the reshape at :262
                        (the RPN blocks after it stay on stock PyTorch in the reference and are
                        out of scope here; `MiddleConvNet.forward` returns the (B,128,H,W) tensor
                        the RPN consumes).

Parameters live in ordinary nn.Linear / nn.BatchNorm1d / nn.Conv3d / nn.BatchNorm3d containers so
that a reference `state_dict()` loads by name (`vfe_1.fcn.0.weight`, `vfe_1.bn.running_mean`,
`middle_layer.0.conv.weight`, `middle_layer.0.batch_norm.weight`, ...), but their forward is never
called: every forward here launches kernels from the C ABI.

BatchNorm: `.eval()` uses the running statistics (folded into the kernels' scale/shift).  `.train()`
(the reference trains with model.train(True), train.py:148) normalises with the BATCH statistics -
over all sum(K)*T rows incl. pad rows for the VFE layers, over B*D*H*W positions for the conv layers -
and updates running_mean / running_var / num_batches_tracked exactly as nn.BatchNorm does
(momentum 0.1, unbiased variance).  In training mode `FeatureLearningNet` and `MiddleConvNet` build an autograd
graph out of three kinds of nodes whose backward launches the library's gradient kernels (`_SVFETrainFn`,
`_Conv1SparseTrainFn`, `_ConvTrainFn`): loss.backward() (train.py:149) fills `.grad` of every hot-path
parameter.  The dense grid itself carries no gradient: the voxel-wise features reach the first conv layer's
node through a side channel (`dense._vn_sparse`), and its data gradient is evaluated at the occupied cells only.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops
from .config import MIDDLE_LAYERS, grid_config


def _fold_bn(bn: nn.modules.batchnorm._BatchNorm):
    """Eval-mode BN as y = x * scale + shift (fp32, computed in fp64 then rounded)."""
    var = bn.running_var.detach().double()
    scale = bn.weight.detach().double() / torch.sqrt(var + bn.eps)
    shift = bn.bias.detach().double() - bn.running_mean.detach().double() * scale
    return scale.float().contiguous(), shift.float().contiguous()


def _fold_batch_stats(bn: nn.modules.batchnorm._BatchNorm, sums: torch.Tensor, count: int):
    """Training-mode BN from per-channel (sum, sum of squares) over `count` elements: returns the folded
    (scale, shift) plus (mean, invstd) for the backward, and updates the running statistics like nn.BatchNorm (momentum, unbiased variance).
    A handful of 16/64-element tensor operations on the device - no host synchronisation."""
    mean = sums[0] / count
    var = (sums[1] / count - mean * mean).clamp_min(0.0)  # biased: what the normalisation uses
    invstd = torch.rsqrt(var + bn.eps)
    scale = bn.weight.detach().double() * invstd
    shift = bn.bias.detach().double() - mean * scale
    with torch.no_grad():
        if bn.track_running_stats and bn.running_mean is not None:
            bn.num_batches_tracked += 1
            m = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
            unbiased = var * (count / max(count - 1, 1))
            bn.running_mean.mul_(1 - m).add_(mean.to(bn.running_mean.dtype), alpha=m)
            bn.running_var.mul_(1 - m).add_(unbiased.to(bn.running_var.dtype), alpha=m)
    return scale.float().contiguous(), shift.float().contiguous(), mean.float().contiguous(), invstd.float().contiguous()


def _require_eval(module: nn.Module) -> None:
    if module.training:
        raise NotImplementedError(
            "%s: this entry point implements eval-mode BatchNorm only; training-mode batch statistics are "
            "implemented by FeatureLearningNet.forward / voxelwise and ConvMD / MiddleConvNet.forward" % type(module).__name__
        )


# ---------------------------------------------------------------------------------------------
# autograd nodes of the training path (backward = the library's gradient kernels)
# ---------------------------------------------------------------------------------------------
class _SVFETrainFn(torch.autograd.Function):
    """feature (K,T,7) -> voxel-wise (K,128) with batch-statistics BatchNorm (model.py:95-100 in train mode)."""

    @staticmethod
    def forward(ctx, feature, w1, b1, g1, be1, w2, b2, g2, be2, net):
        p1, p2 = net.batch_folded(feature)
        bn1, bn2 = net._last_bn
        vw = ops.svfe_fused_fwd(feature, p1, p2)
        ctx.save_for_backward(feature, p1[0], p1[1], p2[0], p2[1], *bn1, *bn2)
        return vw

    @staticmethod
    def backward(ctx, dvw):
        feature, w1, b1, w2, b2, *bn = ctx.saved_tensors
        g = ops.svfe_bwd(feature, dvw.contiguous().float(), w1, b1, w2, b2, bn[0:4], bn[4:8])
        return None, g["w1"], g["b1"], g["gamma1"], g["beta1"], g["w2"], g["b2"], g["gamma2"], g["beta2"], None


def _conv_train_forward(layer, x, final, out, box_flags=None):
    """raw conv (+ batch statistics in the epilogue) -> fold -> BN + ReLU into a separate buffer (the raw
    output is kept for the backward)."""
    packed, _, _ = layer.folded()
    raw, sums = ops.conv3d_raw_stats_fwd(x, packed, layer.conv.bias.detach(), layer.stride[0], layer.padding[0],
                                         box_flags=box_flags)
    B, Dout, H, W, _ = raw.shape
    stats = _fold_batch_stats(layer.batch_norm, sums, B * Dout * H * W)
    y = ops.bn_relu_apply(raw, stats[0], stats[1], final=final, out=out if final else None)
    return y, raw, stats


def _conv_train_backward(x, raw, stats, dy, final, stride_d, pad_d, box_flags=None):
    """-> (draw bf16, dW (64,cin,3,3,3) f32, dgamma, dbeta)."""
    dy = dy.contiguous()
    dy = dy.float() if final else dy.to(torch.bfloat16)
    draw, dgamma, dbeta = ops.bn_relu_bwd(raw, dy, *stats, final=final)
    dw = ops.conv3d_wgrad(x, draw, stride_d, pad_d, box_flags=box_flags)
    return draw, dw, dgamma, dbeta


class _ConvTrainFn(torch.autograd.Function):
    """ConvMD in training mode on bf16 NDHWC (model.py:158-167).  The data gradient exists for the layers with
    64 input channels; the first layer's goes through _Conv1SparseTrainFn."""

    @staticmethod
    def forward(ctx, x, weight, bias, gamma, beta, layer, final):
        y, raw, stats = _conv_train_forward(layer, x, final, None)
        ctx.save_for_backward(x, raw, weight, *stats)
        ctx.geom = (layer.stride[0], layer.padding[0], final)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, raw, weight, *stats = ctx.saved_tensors
        stride_d, pad_d, final = ctx.geom
        draw, dw, dgamma, dbeta = _conv_train_backward(x, raw, stats, dy, final, stride_d, pad_d)
        dx = None
        if ctx.needs_input_grad[0]:
            if x.shape[-1] != 64:
                raise RuntimeError("the first middle layer has no dense data gradient; feed it FeatureLearningNet's output")
            dx = ops.conv3d_dgrad(draw, ops.conv3d_pack_weights_dgrad(weight), x.shape[1], stride_d, pad_d)
        # the conv bias shifts the batch mean only: its gradient under batch-statistics BN is exactly zero
        return dx, dw, torch.zeros_like(dgamma), dgamma, dbeta, None, None


class _Conv1SparseTrainFn(torch.autograd.Function):
    """to_dense (model.py:102-106) + first ConvMD in training mode.  `dense` already holds the scattered bf16
    voxel-wise rows; the gradient flows to `vw` (K,128) through the occupied cells only."""

    @staticmethod
    def forward(ctx, vw, coords, dense, weight, bias, gamma, beta, layer):
        # boxes of the grid no voxel lies in are all-zero: their MMAs are skipped in the forward and in the
        # weight gradient (exact); every output block is still written
        B, D, H, W, _ = dense.shape
        flags = ops.conv3d_new_box_flags(B, D, H, W, dense.device)
        ops.conv3d_mark_boxes(coords, B, D, H, W, [(layer.stride[0], layer.padding[0])], [flags])
        y, raw, stats = _conv_train_forward(layer, dense, False, None, box_flags=flags)
        ctx.save_for_backward(dense, raw, weight, coords, flags, *stats)
        ctx.geom = (layer.stride[0], layer.padding[0])
        return y

    @staticmethod
    def backward(ctx, dy):
        dense, raw, weight, coords, flags, *stats = ctx.saved_tensors
        stride_d, pad_d = ctx.geom
        draw, dw, dgamma, dbeta = _conv_train_backward(dense, raw, stats, dy, False, stride_d, pad_d, box_flags=flags)
        dvw = None
        if ctx.needs_input_grad[0]:
            dvw = ops.conv1_dgrad_gather(draw, ops.conv1_weight_taps(weight), coords, stride_d, pad_d)
        return dvw, None, None, dw, torch.zeros_like(dgamma), dgamma, dbeta, None


class _VFELayerTrainFn(torch.autograd.Function):
    """A stand-alone VFELayer in training mode (model.py:74-82).  Forward: the library kernels (statistics kernel +
    vfe_layer_fwd with the batch statistics folded).  Backward: a convenience outside the hot path - training goes
    through FeatureLearningNet, whose backward is the fused `vn_svfe_bwd` kernels - so it is written with tensor
    operations (recompute Linear -> ReLU -> batch-statistics BN -> max -> concat -> mask and differentiate that)."""

    @staticmethod
    def forward(ctx, inputs, mask, weight, bias, gamma, beta, stats, eps):
        ctx.save_for_backward(inputs, mask, weight, bias, gamma, beta)
        ctx.eps = eps
        return ops.vfe_layer_fwd(inputs, mask, weight.detach().contiguous(), bias.detach().contiguous(), stats[0], stats[1])

    @staticmethod
    def backward(ctx, dout):
        inputs, mask, weight, bias, gamma, beta = ctx.saved_tensors
        with torch.enable_grad():
            x = inputs.detach().requires_grad_(True)
            leaves = [t.detach().requires_grad_(True) for t in (weight, bias, gamma, beta)]
            w, b, g, be = leaves
            pw = torch.relu(torch.nn.functional.linear(x, w, b))
            flat = pw.reshape(-1, pw.shape[-1])
            mean, var = flat.mean(0), flat.var(0, unbiased=False)
            pw = (pw - mean) * torch.rsqrt(var + ctx.eps) * g + be
            agg = pw.max(dim=1, keepdim=True)[0]
            out = torch.cat([pw, agg.expand(-1, pw.shape[1], -1)], dim=2) * mask.reshape(pw.shape[0], pw.shape[1], 1).float()
            grads = torch.autograd.grad(out, [x] + leaves, dout)
        return grads[0], None, grads[1], grads[2], grads[3], grads[4], None, None


class VFELayer(nn.Module):
    """model.py:60-82.  forward(inputs (K,T,cin) f32, mask (K,T,1) bool) -> (K,T,cout) f32."""

    def __init__(self, cin: int, cout: int):
        super().__init__()
        self.in_channels = cin
        self.out_channels = cout
        self.local_agg_features = cout // 2
        self.fcn = nn.Sequential(nn.Linear(self.in_channels, self.local_agg_features), nn.ReLU())
        self.bn = nn.BatchNorm1d(self.local_agg_features)

    def folded(self):
        """(weight, bias, bn_scale, bn_shift): Linear -> ReLU -> BN as relu(Wx+b)*scale+shift."""
        scale, shift = _fold_bn(self.bn)
        lin = self.fcn[0]
        return lin.weight.detach().contiguous(), lin.bias.detach().contiguous(), scale, shift

    def forward(self, inputs, mask):
        inputs = inputs.contiguous()
        if not self.training:
            return ops.vfe_layer_fwd(inputs, mask, *self.folded())
        # model.py:74-76 under .train(): BatchNorm1d with the statistics of THIS input - all K*T rows of
        # relu(Linear(x)), pad rows included - and the usual running-statistics update.
        lin = self.fcn[0]
        w, b = lin.weight.detach().contiguous(), lin.bias.detach().contiguous()
        K, T, _ = inputs.shape
        if K == 0:
            raise ValueError("Expected more than 1 value per channel when training (an empty voxel batch)")
        stats = _fold_batch_stats(self.bn, ops.vfe_layer_batch_stats(inputs.detach(), w, b), K * T)
        needs_grad = torch.is_grad_enabled() and (inputs.requires_grad or any(p.requires_grad for p in self.parameters()))
        if needs_grad:
            return _VFELayerTrainFn.apply(inputs, mask, lin.weight, lin.bias, self.bn.weight, self.bn.bias, stats, self.bn.eps)
        return ops.vfe_layer_fwd(inputs, mask, w, b, stats[0], stats[1])


class FeatureLearningNet(nn.Module):
    """model.py:85-108.  forward(feature: list[(K_i,T,7) f32], coordinate: list[(K_i,4) i64])
    -> dense (B, D, H, W, 128) f32, B = len(feature)."""

    def __init__(self, target: str = "Car"):
        super().__init__()
        self.target = target
        self.vfe_1 = VFELayer(7, 32)
        self.vfe_2 = VFELayer(32, 128)
        # dtype of the dense grid `forward` returns: float32 is the reference's contract (model.py:102-106);
        # bfloat16 is what the middle layers consume and saves the 2x larger buffer and its conversion
        self.dense_dtype = torch.float32

    def batch_folded(self, feature: torch.Tensor):
        """Training mode: both VFE BatchNorms folded from the statistics of THIS batch (all K*T rows, pads
        included - SURVEY F6); updates the running statistics.  Two reduction passes: VFE-2's input depends
        on VFE-1's normalised output."""
        K, T, _ = feature.shape
        l1, l2 = self.vfe_1.fcn[0], self.vfe_2.fcn[0]
        w1, b1 = l1.weight.detach().contiguous(), l1.bias.detach().contiguous()
        w2, b2 = l2.weight.detach().contiguous(), l2.bias.detach().contiguous()
        bn1 = _fold_batch_stats(self.vfe_1.bn, ops.svfe_batch_stats(feature, 1, w1, b1), K * T)
        s1, t1 = bn1[0], bn1[1]
        bn2 = _fold_batch_stats(self.vfe_2.bn, ops.svfe_batch_stats(feature, 2, w1, b1, s1, t1, w2, b2), K * T)
        self._last_bn = (bn1, bn2)  # (scale, shift, mean, invstd) of this batch, read by the backward node
        return (w1, b1, s1, t1), (w2, b2, bn2[0], bn2[1])

    def voxelwise(self, feature: torch.Tensor, out=None, n_voxels_dev=None) -> torch.Tensor:
        """model.py:95-100 as one fused kernel: (K,T,7) -> (K,128)."""
        if self.training:
            if n_voxels_dev is not None:
                raise NotImplementedError("training-mode statistics need the exact K (pass feature[:K])")
            l1, l2 = self.vfe_1, self.vfe_2
            return _SVFETrainFn.apply(feature, l1.fcn[0].weight, l1.fcn[0].bias, l1.bn.weight, l1.bn.bias,
                                      l2.fcn[0].weight, l2.fcn[0].bias, l2.bn.weight, l2.bn.bias, self)
        else:
            p1, p2 = self.vfe_1.folded(), self.vfe_2.folded()
        return ops.svfe_fused_fwd(feature, p1, p2, out=out, n_voxels_dev=n_voxels_dev)

    def forward(self, feature: list, coordinate: list):
        bs = len(feature)
        feature = torch.cat(feature, dim=0).contiguous()
        coordinate = torch.cat(coordinate, dim=0).contiguous()
        vw = self.voxelwise(feature)
        g = grid_config(self.target)
        dense = torch.empty((bs, g.depth, g.height, g.width, 128), dtype=self.dense_dtype, device=vw.device)
        ops.fill_zero(dense)
        if self.dense_dtype == torch.bfloat16:
            ops.scatter_dense_bf16(vw.detach(), coordinate, dense)
        else:
            ops.scatter_dense_f32(vw.detach(), coordinate, dense)
        if vw.requires_grad:  # the gradient bypasses the dense grid: MiddleConvNet picks this up
            dense._vn_sparse = (vw, coordinate)
        return dense


class ConvMD(nn.Module):
    """model.py:111-167, 3-D case: Conv3d(k=3) -> BatchNorm3d -> ReLU.

    forward(x) takes the logical NCDHW fp32 tensor the reference passes (model.py:259 hands it a
    channels-last view) and returns logical NCDHW fp32.  Internally: bf16 NDHWC in, tcgen05
    implicit GEMM with fp32 accumulation, BN/bias/ReLU in the epilogue.  `forward_ndhwc` is the
    chaining form without layout or dtype round trips.
    """

    def __init__(self, input_dim: int, cin: int, cout: int, kernel_size: int, stride, padding, bn: bool = True,
                 activation: bool = True):
        super().__init__()
        if input_dim != 3:
            raise ValueError("the B200 path implements the 3-D middle layers; 2-D ConvMD (the RPN) stays on stock PyTorch")
        if kernel_size != 3 or cout != 64 or cin not in (64, 128) or not bn or not activation:
            raise ValueError("unsupported ConvMD configuration for the B200 kernel (model.py:207-209 shapes only)")
        stride, padding = tuple(stride), tuple(padding)
        if stride[1:] != (1, 1) or padding[1:] != (1, 1) or (stride[0], padding[0]) not in ((2, 1), (1, 0)):
            raise ValueError("unsupported stride/padding %s/%s" % (stride, padding))
        self.input_dim, self.cin, self.cout = input_dim, cin, cout
        self.kernel_size, self.stride, self.padding = kernel_size, stride, padding
        self.bn, self.activation = bn, activation
        self.conv = nn.Conv3d(cin, cout, kernel_size, stride, padding)
        self.batch_norm = nn.BatchNorm3d(cout)
        self._packed = None
        self._packed_key = None

    def folded(self):
        """(packed bf16 weights, scale, shift): BN(conv(x)+b) = conv(x)*scale + shift."""
        w = self.conv.weight
        key = (w.data_ptr(), w._version, w.device)
        if self._packed is None or self._packed_key != key:
            self._packed = ops.conv3d_pack_weights(w, self.stride[0])
            self._packed_key = key
        scale, bn_shift = _fold_bn(self.batch_norm)
        shift = (self.conv.bias.detach().double() * scale.double() + bn_shift.double()).float().contiguous()
        return self._packed, scale, shift

    def forward_ndhwc(self, x_bf16, final: bool = False, out=None, sparse=None):
        if self.training:  # model.py:159-165 with batch statistics: raw conv (+ sums) -> fold -> BN + ReLU
            if torch.is_grad_enabled():
                if sparse is not None:
                    vw, coords = sparse
                    return _Conv1SparseTrainFn.apply(vw, coords, x_bf16, self.conv.weight, self.conv.bias,
                                                     self.batch_norm.weight, self.batch_norm.bias, self)
                # (a caller-provided `out` is not used here: the node owns its output)
                return _ConvTrainFn.apply(x_bf16, self.conv.weight, self.conv.bias, self.batch_norm.weight,
                                          self.batch_norm.bias, self, final)
            packed, _, _ = self.folded()
            raw, sums = ops.conv3d_raw_stats_fwd(x_bf16, packed, self.conv.bias.detach(), self.stride[0], self.padding[0])
            B, Dout, H, W, _ = raw.shape
            scale, shift, _, _ = _fold_batch_stats(self.batch_norm, sums, B * Dout * H * W)
            return ops.bn_relu_apply(raw, scale, shift, final=final, out=out if final else raw)
        packed, scale, shift = self.folded()
        return ops.conv3d_bn_relu_fwd(x_bf16, packed, scale, shift, self.stride[0], self.padding[0], final=final, out=out)

    def forward(self, x):
        x_ndhwc = x.permute(0, 2, 3, 4, 1).contiguous().to(torch.bfloat16)
        y = self.forward_ndhwc(x_ndhwc)
        return y.permute(0, 4, 1, 2, 3).float()


class MiddleConvNet(nn.Module):
    """model.py:202-262 up to the reshape: forward(x (B,D,H,W,128) f32 or bf16) -> (B,128,H,W) f32."""

    def __init__(self):
        super().__init__()
        self.middle_layer = nn.Sequential(
            *[ConvMD(3, l["cin"], l["cout"], 3, l["stride"], l["padding"]) for l in MIDDLE_LAYERS]
        )

    def forward(self, x, out=None):
        sparse = getattr(x, "_vn_sparse", None) if self.training else None
        if x.dtype != torch.bfloat16:
            x = x.to(torch.bfloat16)
        x = x.contiguous()
        x = self.middle_layer[0].forward_ndhwc(x, sparse=sparse)
        x = self.middle_layer[1].forward_ndhwc(x)
        # the last layer writes fp32 straight into the (B, c*2+d, H, W) layout of model.py:262
        return self.middle_layer[2].forward_ndhwc(x, final=True, out=out)
