"""TEST INFRASTRUCTURE - CPU fp32 oracle for the stacked VFE, the sparse->dense scatter and the
Conv3d middle layers.

Functional torch-CPU restatement (fp32, no modules) of

* `VFELayer.forward`            voxelnet/model.py:74-82
* `FeatureLearningNet.forward`  voxelnet/model.py:91-108
* `ConvMD.forward` (3-D)        voxelnet/model.py:158-167
* `MiddleConvNet.forward` head  voxelnet/model.py:257-262 (permute, middle_layer, reshape)

driven by plain state dicts that use the reference's parameter names
(`vfe_1.fcn.0.weight`, `vfe_1.bn.running_mean`, `middle_layer.0.conv.weight`, ...), so a
`state_dict()` taken from the reference modules can be fed in unchanged.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu-baseline / `--impl reference` legs may
import this module.  Parity pinning: see `oracle/voxelize_np.py` - the reference holds no golden
vectors for this path; `tests/test_oracle_vs_reference.py` checks these functions against the
imported reference modules (container only) and `tests/golden/` freezes reference outputs.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-5  # torch default used by nn.BatchNorm1d/3d in model.py:72,153

# (cin, cout, stride, padding) of the three ConvMD(3, ...) at model.py:207-209
MIDDLE_LAYERS = (
    (128, 64, (2, 1, 1), (1, 1, 1)),
    (64, 64, (1, 1, 1), (0, 1, 1)),
    (64, 64, (2, 1, 1), (1, 1, 1)),
)


def grid_dims(target: str):
    """(D, H, W, T) - voxelnet/config.py:38-50 (Car) and :61-73 (Pedestrian/Cyclist)."""
    return (10, 400, 352, 35) if target == "Car" else (10, 200, 240, 45)


def _bn_train(x, sd, prefix, new_stats):
    """nn.BatchNorm in training mode (the reference trains with model.train(True), train.py:148): batch
    statistics, momentum 0.1; the updated running statistics are written to `new_stats` (if a dict)."""
    rm, rv = sd[prefix + "running_mean"].clone(), sd[prefix + "running_var"].clone()
    y = F.batch_norm(x, rm, rv, sd[prefix + "weight"], sd[prefix + "bias"], training=True, momentum=0.1, eps=BN_EPS)
    if new_stats is not None:
        new_stats[prefix + "running_mean"], new_stats[prefix + "running_var"] = rm, rv
    return y


def _bn_eval(x, sd, prefix, training=False, new_stats=None):
    if training:
        return _bn_train(x, sd, prefix, new_stats)
    return F.batch_norm(
        x,
        sd[prefix + "running_mean"],
        sd[prefix + "running_var"],
        sd[prefix + "weight"],
        sd[prefix + "bias"],
        training=False,
        eps=BN_EPS,
    )


def vfe_layer(inputs: torch.Tensor, mask: torch.Tensor, sd: dict, prefix: str, training=False, new_stats=None) -> torch.Tensor:
    """model.py:74-82.  inputs (K,T,cin) f32, mask (K,T,1) bool -> (K,T,cout) f32.

    Linear -> ReLU -> BatchNorm1d over channels (eval mode) -> max over ALL T rows (unmasked)
    -> concat [pointwise, repeated max] -> multiply by mask.
    """
    T = inputs.shape[1]
    pw = F.relu(F.linear(inputs, sd[prefix + "fcn.0.weight"], sd[prefix + "fcn.0.bias"]))
    pw = _bn_eval(pw.transpose(1, 2), sd, prefix + "bn.", training, new_stats).transpose(1, 2)
    agg = pw.max(dim=1, keepdim=True)[0]
    cat = torch.cat([pw, agg.expand(-1, T, -1)], dim=2)
    return cat * mask.expand(-1, -1, cat.shape[2]).float()


def feature_mask(feature: torch.Tensor) -> torch.Tensor:
    """model.py:95-96: a row is 'on' iff the max over its 7 features is != 0."""
    return feature.max(dim=2, keepdim=True)[0] != 0


def svfe_voxelwise(feature: torch.Tensor, sd: dict, training=False, new_stats=None) -> torch.Tensor:
    """model.py:95-100: (sum K, T, 7) -> (sum K, 128)."""
    mask = feature_mask(feature)
    x = vfe_layer(feature, mask, sd, "vfe_1.", training, new_stats)
    x = vfe_layer(x, mask, sd, "vfe_2.", training, new_stats)
    return x.max(dim=1)[0]


def scatter_dense(voxelwise: torch.Tensor, coordinate: torch.Tensor, batch: int, target: str) -> torch.Tensor:
    """model.py:102-106: COO (b,z,y,x) -> dense (B,D,H,W,128) f32 (duplicates would be summed)."""
    D, H, W, _ = grid_dims(target)
    dense = torch.zeros((batch, D, H, W, voxelwise.shape[1]), dtype=voxelwise.dtype)
    c = coordinate.long()
    dense.index_put_((c[:, 0], c[:, 1], c[:, 2], c[:, 3]), voxelwise, accumulate=True)
    return dense


def feature_learning_net(feature: list, coordinate: list, sd: dict, target: str = "Car") -> torch.Tensor:
    """model.py:91-108 with list inputs as produced by collate_fn (dataset.py:88-96)."""
    vw = svfe_voxelwise(torch.cat(feature, dim=0), sd)
    return scatter_dense(vw, torch.cat(coordinate, dim=0), len(feature), target)


def conv_md3(x: torch.Tensor, sd: dict, prefix: str, stride, padding, training=False, new_stats=None) -> torch.Tensor:
    """model.py:158-167 for input_dim == 3: Conv3d(k=3) -> BatchNorm3d (eval) -> ReLU."""
    x = F.conv3d(x, sd[prefix + "conv.weight"], sd[prefix + "conv.bias"], stride=stride, padding=padding)
    x = _bn_eval(x, sd, prefix + "batch_norm.", training, new_stats)
    return F.relu(x)


def middle_layers(dense: torch.Tensor, sd: dict, upto: int = 3, prefix: str = "middle_layer.", training=False,
                  new_stats=None) -> torch.Tensor:
    """model.py:257-262: (B,D,H,W,C) -> permute -> 3 x ConvMD -> reshape (B,128,H,W).

    `upto` < 3 returns the NCDHW activation after that many layers (for per-layer checks).
    """
    B, _, H, W, _ = dense.shape
    x = dense.permute(0, 4, 1, 2, 3)
    for i, (_, _, stride, padding) in enumerate(MIDDLE_LAYERS[:upto]):
        x = conv_md3(x, sd, "%s%d." % (prefix, i), stride, padding, training, new_stats)
    if upto < 3:
        return x
    return x.reshape((B, -1, H, W))


def random_state_dict(seed: int = 0, nontrivial_bn: bool = True) -> dict:
    """Reference-shaped parameters (names as in SURVEY section 5 'Checkpoint'): default torch init for
    Linear/Conv3d, BatchNorm affine and running statistics randomised so that BN is not the
    identity.  Keys: vfe_{1,2}.{fcn.0,bn}.*, middle_layer.{0,1,2}.{conv,batch_norm}.*
    """
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def uniform(shape, bound):
        return (torch.rand(shape, generator=g) * 2 - 1) * bound

    def bn(prefix, c):
        if nontrivial_bn:
            sd[prefix + "weight"] = 0.5 + torch.rand(c, generator=g)
            sd[prefix + "bias"] = uniform((c,), 0.3)
            sd[prefix + "running_mean"] = uniform((c,), 0.2)
            sd[prefix + "running_var"] = 0.5 + torch.rand(c, generator=g)
        else:
            sd[prefix + "weight"] = torch.ones(c)
            sd[prefix + "bias"] = torch.zeros(c)
            sd[prefix + "running_mean"] = torch.zeros(c)
            sd[prefix + "running_var"] = torch.ones(c)
        sd[prefix + "num_batches_tracked"] = torch.zeros((), dtype=torch.long)

    for name, cin, cout in (("vfe_1.", 7, 32), ("vfe_2.", 32, 128)):
        units = cout // 2
        bound = 1.0 / cin ** 0.5  # nn.Linear default (kaiming_uniform a=sqrt(5))
        sd[name + "fcn.0.weight"] = uniform((units, cin), bound)
        sd[name + "fcn.0.bias"] = uniform((units,), bound)
        bn(name + "bn.", units)
    for i, (cin, cout, _, _) in enumerate(MIDDLE_LAYERS):
        bound = 1.0 / (cin * 27) ** 0.5
        sd["middle_layer.%d.conv.weight" % i] = uniform((cout, cin, 3, 3, 3), bound)
        sd["middle_layer.%d.conv.bias" % i] = uniform((cout,), bound)
        bn("middle_layer.%d.batch_norm." % i, cout)
    return sd


# ---------------------------------------------------------------------------------------------
# bf16-emulating variant of the middle layers (checker for the GRADIENTS of the bf16 CUDA path)
# ---------------------------------------------------------------------------------------------
class _RoundBF16(torch.autograd.Function):
    """x -> bf16 -> f32 with a straight-through gradient."""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return g


def middle_layers_bf16_emulated(dense: torch.Tensor, sd: dict, prefix: str = "middle_layer.", return_acts: bool = False):
    """model.py:257-262 in training mode, fp32 arithmetic on values rounded to bf16 at the points where the CUDA
    path stores bf16 (layer inputs, weights, the raw convolution output, the two intermediate activations).

    Why it exists: against the plain fp32 oracle the forward differs by ~1e-2 of max, which flips the ReLU of the
    few percent of elements that sit that close to zero; with a white-noise upstream gradient every flipped
    element moves a gradient sum by O(1), so gradient errors of ~1e-1 of max against plain fp32 are rounding
    noise, not bugs.  With the rounding points emulated the masks agree and the comparison is tight."""
    rnd = _RoundBF16.apply
    B, _, H, W, _ = dense.shape
    x = rnd(dense).permute(0, 4, 1, 2, 3)
    acts = []
    for i, (_, _, stride, padding) in enumerate(MIDDLE_LAYERS):
        p = "%s%d." % (prefix, i)
        raw32 = F.conv3d(x, rnd(sd[p + "conv.weight"]), sd[p + "conv.bias"], stride=stride, padding=padding)
        mean = raw32.mean(dim=(0, 2, 3, 4), keepdim=True)
        var = raw32.var(dim=(0, 2, 3, 4), unbiased=False, keepdim=True)
        scale = sd[p + "batch_norm.weight"].view(1, -1, 1, 1, 1) * torch.rsqrt(var + BN_EPS)
        shift = sd[p + "batch_norm.bias"].view(1, -1, 1, 1, 1) - mean * scale
        x = F.relu(rnd(raw32) * scale + shift)
        if i < 2:
            x = rnd(x)
        if return_acts:
            x.retain_grad()
            acts.append(x)
    out = x.reshape((B, -1, H, W))
    return (out, acts) if return_acts else out


def fill_rpn_parameters(module, seed: int = 0) -> str:
    """Deterministic, NAME-keyed values for every RPN parameter and buffer of `module` (the reference's `MiddleConvNet`,
    or any mirror with the same names: block1..3, deconv1..3, prob_conv, reg_conv; voxelnet/model.py:212-254), so that the
    reference in the container and a mirror on the GPU box hold the same numbers without shipping 18 MB of weights.
    Weights: uniform with the He bound of their fan-in (activations keep their scale through the 20 layers); BatchNorm
    scale 0.5..1.5, shift +-0.2, running mean +-0.15, running variance 0.5..1.5.  `middle_layer.*` is left alone.
    Returns the sha256 of the filled values (the golden stores it)."""
    import hashlib
    import zlib

    h = hashlib.sha256()
    with torch.no_grad():
        for name, t in module.state_dict().items():
            if name.startswith("middle_layer.") or name.endswith("num_batches_tracked"):
                continue
            g = torch.Generator().manual_seed((zlib.crc32(name.encode()) ^ (seed * 2654435761)) & 0x7FFFFFFF)
            u = torch.rand(t.shape, generator=g, dtype=torch.float32)
            if name.endswith("running_var"):
                v = 0.5 + u
            elif name.endswith("running_mean"):
                v = (u - 0.5) * 0.3
            elif name.endswith("batch_norm.weight"):
                v = 0.5 + u
            elif name.endswith("batch_norm.bias"):
                v = (u - 0.5) * 0.4
            elif name.endswith("deconv.weight"):  # (cin, cout, k, k), kernel = stride except the 3x3 stride-1 deconv1
                k = t.shape[2]
                taps = 9 if k == 3 else 1
                v = (u - 0.5) * 2.0 * (6.0 / (t.shape[0] * taps)) ** 0.5
            elif name.endswith(".weight"):        # (cout, cin, k, k)
                v = (u - 0.5) * 2.0 * (6.0 / (t.shape[1] * t.shape[2] * t.shape[3])) ** 0.5
            else:                                  # conv / deconv bias
                v = (u - 0.5) * 0.2
            t.copy_(v.to(t.dtype))
            h.update(name.encode())
            h.update(v.contiguous().numpy().tobytes())
    return h.hexdigest()


def rpn_golden_input(seed: int = 0, shape=(1, 128, 64, 48)) -> torch.Tensor:
    """The seeded non-negative map the RPN golden is computed on (what the ReLU of the last 3-D layer hands over)."""
    g = torch.Generator().manual_seed(1000 + seed)
    return torch.randn(shape, generator=g).relu()
