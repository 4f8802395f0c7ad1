"""Thin torch-tensor wrappers over the C ABI (include/voxelnet_b200.h).

torch is used here for device memory and the current stream only; every arithmetic step is a
kernel of libvoxelnet_b200.so.  All functions require CUDA tensors and raise otherwise - there is
no CPU path.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _cabi
from ._cabi import check


def _require_cuda(*tensors) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError(
                "voxelnet_pytorch_b200 runs on CUDA only (sm_100a kernels, no CPU fallback); got a %s tensor" % t.device
            )


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _contig(t, dtype):
    if t.dtype != dtype or not t.is_contiguous():
        raise RuntimeError("expected a contiguous %s tensor, got %s contiguous=%s" % (dtype, t.dtype, t.is_contiguous()))
    return t


# ---------------------------------------------------------------------------------------------
# voxelizer
# ---------------------------------------------------------------------------------------------
def voxelize_workspace_bytes(grid, n_frames: int, n_points: int) -> int:
    return int(_cabi.lib().vn_voxelize_workspace_bytes(C.byref(grid), n_frames, n_points))


def voxelize_index(grid, points, frame_offsets, max_points_per_frame: int, voxel_offsets, workspace) -> None:
    """Phase 1 (see header).  points (N,4) f32, frame_offsets (B+1) i32, voxel_offsets (B+1) i32 out."""
    _require_cuda(points, frame_offsets, voxel_offsets, workspace)
    _contig(points, torch.float32), _contig(frame_offsets, torch.int32), _contig(voxel_offsets, torch.int32)
    n_frames = frame_offsets.numel() - 1
    with torch.cuda.device(points.device):
        check(
            _cabi.lib().vn_voxelize_index(
                C.byref(grid), _ptr(points), _ptr(frame_offsets), n_frames, points.shape[0], max_points_per_frame,
                _ptr(voxel_offsets), _ptr(workspace), workspace.numel() * workspace.element_size(), _stream(),
            ),
            "vn_voxelize_index",
        )


def voxelize_fill(grid, points, frame_offsets, max_points_per_frame, voxel_offsets, feature, coords, counts, workspace) -> None:
    """Phase 2.  feature (cap,T,7) f32, coords (cap,4) i64, counts (cap,) i64 out."""
    _require_cuda(points, frame_offsets, voxel_offsets, feature, coords, counts, workspace)
    _contig(feature, torch.float32), _contig(coords, torch.int64), _contig(counts, torch.int64)
    n_frames = frame_offsets.numel() - 1
    with torch.cuda.device(points.device):
        check(
            _cabi.lib().vn_voxelize_fill(
                C.byref(grid), _ptr(points), _ptr(frame_offsets), n_frames, points.shape[0], max_points_per_frame,
                _ptr(voxel_offsets), feature.shape[0], _ptr(feature), _ptr(coords), _ptr(counts), _ptr(workspace),
                workspace.numel() * workspace.element_size(), _stream(),
            ),
            "vn_voxelize_fill",
        )


def voxel_offsets_to_host(voxel_offsets) -> list:
    """Blocking read of the (B+1) voxel prefix (the only synchronising entry point)."""
    _require_cuda(voxel_offsets)
    n = voxel_offsets.numel()
    host = (C.c_int32 * n)()
    with torch.cuda.device(voxel_offsets.device):
        check(_cabi.lib().vn_voxelize_counts_host(_ptr(voxel_offsets), n - 1, host, _stream()), "vn_voxelize_counts_host")
    return list(host)


# ---------------------------------------------------------------------------------------------
# VFE
# ---------------------------------------------------------------------------------------------
def feature_mask(feature) -> torch.Tensor:
    """model.py:95-96 -> (K,T,1) bool."""
    _require_cuda(feature)
    _contig(feature, torch.float32)
    K, T, cin = feature.shape
    mask = torch.empty((K, T, 1), dtype=torch.bool, device=feature.device)
    with torch.cuda.device(feature.device):
        check(_cabi.lib().vn_feature_mask(_ptr(feature), K, T, cin, _ptr(mask), _stream()), "vn_feature_mask")
    return mask


def vfe_layer_fwd(inputs, mask, weight, bias, bn_scale, bn_shift) -> torch.Tensor:
    """model.py:74-82 with eval BN folded.  inputs (K,T,cin) f32, mask (K,T,1) bool -> (K,T,2*units)."""
    _require_cuda(inputs, mask, weight, bias, bn_scale, bn_shift)
    _contig(inputs, torch.float32)
    K, T, cin = inputs.shape
    units = weight.shape[0]
    if mask.dtype != torch.bool or mask.numel() != K * T:
        raise RuntimeError("mask must be a bool tensor with K*T elements")
    mask = mask.contiguous()
    out = torch.empty((K, T, 2 * units), dtype=torch.float32, device=inputs.device)
    with torch.cuda.device(inputs.device):
        check(
            _cabi.lib().vn_vfe_layer_fwd(
                _ptr(inputs), _ptr(mask), K, T, cin, units, _ptr(_contig(weight, torch.float32)),
                _ptr(_contig(bias, torch.float32)), _ptr(_contig(bn_scale, torch.float32)),
                _ptr(_contig(bn_shift, torch.float32)), _ptr(out), _stream(),
            ),
            "vn_vfe_layer_fwd",
        )
    return out


def svfe_fused_fwd(feature, p1, p2, out=None, n_voxels_dev=None) -> torch.Tensor:
    """model.py:95-100 fused.  p1/p2 = (weight, bias, bn_scale, bn_shift) of VFE-1 / VFE-2."""
    _require_cuda(feature, *p1, *p2)
    _contig(feature, torch.float32)
    K, T, cin = feature.shape
    if cin != 7:
        raise RuntimeError("svfe_fused_fwd expects 7 input features per point")
    if out is None:
        out = torch.empty((K, 128), dtype=torch.float32, device=feature.device)
    with torch.cuda.device(feature.device):
        check(
            _cabi.lib().vn_svfe_fused_fwd(
                _ptr(feature), K, T, _ptr(n_voxels_dev), *[_ptr(_contig(t, torch.float32)) for t in p1],
                *[_ptr(_contig(t, torch.float32)) for t in p2], _ptr(out), _stream(),
            ),
            "vn_svfe_fused_fwd",
        )
    return out


def voxelize_bucket_coords(grid, points, frame_offsets, voxel_offsets, coords, workspace) -> None:
    """Bucketing step of voxelizer phase 2 + the (cap,4) i64 voxel coordinates, ahead of voxelize_svfe_fused(...,
    bucketed=True): what depends on the coordinates only can then run next to the SVFE kernel."""
    _require_cuda(points, frame_offsets, voxel_offsets, coords, workspace)
    _contig(coords, torch.int64)
    n_frames = frame_offsets.numel() - 1
    with torch.cuda.device(points.device):
        check(
            _cabi.lib().vn_voxelize_bucket_coords(C.byref(grid), _ptr(points), _ptr(frame_offsets), n_frames, points.shape[0],
                                                  _ptr(voxel_offsets), coords.shape[0], _ptr(coords), _ptr(workspace),
                                                  workspace.numel() * workspace.element_size(), _stream()),
            "vn_voxelize_bucket_coords",
        )


def voxelize_svfe_fused(grid, points, frame_offsets, voxel_offsets, p1, p2, voxelwise, coords, counts, workspace,
                        dense=None, bucketed: bool = False) -> None:
    """Voxelizer phase 2 fused with the stacked VFE: writes coords (cap,4) i64 (counts optional) and the
    voxel-wise maxima - as voxelwise (cap,128) f32 and/or straight into the resident bf16 grid
    `dense` (B,D,H,W,128) - without materialising the (K,T,7) feature buffer."""
    _require_cuda(points, frame_offsets, voxel_offsets, voxelwise, coords, counts, workspace, dense, *p1, *p2)
    _contig(coords, torch.int64)
    if voxelwise is None and dense is None:
        raise RuntimeError("voxelize_svfe_fused needs voxelwise and/or dense")
    if voxelwise is not None:
        _contig(voxelwise, torch.float32)
        if voxelwise.shape[0] != coords.shape[0]:
            raise RuntimeError("voxelwise and coords capacities differ")
    n_frames = frame_offsets.numel() - 1
    if dense is not None:
        _contig(dense, torch.bfloat16)
        if tuple(dense.shape) != (n_frames, grid.D, grid.H, grid.W, 128):
            raise RuntimeError("dense must be (B,D,H,W,128) bf16 for this grid; got %s" % (tuple(dense.shape),))
    fn = _cabi.lib().vn_voxelize_svfe_bucketed if bucketed else _cabi.lib().vn_voxelize_svfe_fused
    with torch.cuda.device(points.device):
        check(
            fn(
                C.byref(grid), _ptr(points), _ptr(frame_offsets), n_frames, points.shape[0], _ptr(voxel_offsets),
                coords.shape[0], *[_ptr(_contig(t, torch.float32)) for t in p1],
                *[_ptr(_contig(t, torch.float32)) for t in p2], _ptr(voxelwise), _ptr(dense), _ptr(coords), _ptr(counts),
                _ptr(workspace), workspace.numel() * workspace.element_size(), _stream(),
            ),
            "vn_voxelize_svfe_fused",
        )


def voxelize_svfe_rows(grid, points, frame_offsets, voxel_offsets, p1, p2, rows, coords, workspace, bucketed: bool = False) -> None:
    """voxelize_svfe_fused without any dense tensor (model.py:102-106 never materialised): the voxel-wise maxima of voxel
    row r go to rows[r] ((cap,128) bf16, np.unique row order); the cell -> row map is the occupancy index the voxelizer
    left in `workspace` (consumed by conv3d_bn_relu_fwd_rows)."""
    _require_cuda(points, frame_offsets, voxel_offsets, rows, coords, workspace, *p1, *p2)
    _contig(coords, torch.int64), _contig(rows, torch.bfloat16)
    if rows.dim() != 2 or rows.shape[1] != 128 or rows.shape[0] != coords.shape[0]:
        raise RuntimeError("rows must be (cap,128) bf16 with the capacity of coords; got %s" % (tuple(rows.shape),))
    n_frames = frame_offsets.numel() - 1
    with torch.cuda.device(points.device):
        check(
            _cabi.lib().vn_voxelize_svfe_rows(
                C.byref(grid), _ptr(points), _ptr(frame_offsets), n_frames, points.shape[0], _ptr(voxel_offsets),
                coords.shape[0], *[_ptr(_contig(t, torch.float32)) for t in p1],
                *[_ptr(_contig(t, torch.float32)) for t in p2], _ptr(rows), _ptr(coords), 1 if bucketed else 0,
                _ptr(workspace), workspace.numel() * workspace.element_size(), _stream(),
            ),
            "vn_voxelize_svfe_rows",
        )


# ---------------------------------------------------------------------------------------------
# sparse -> dense
# ---------------------------------------------------------------------------------------------
def fill_zero(t) -> None:
    _require_cuda(t)
    with torch.cuda.device(t.device):
        check(_cabi.lib().vn_fill_zero(_ptr(t), t.numel() * t.element_size(), _stream()), "vn_fill_zero")


def scatter_dense_f32(voxelwise, coords, dense, n_voxels_dev=None) -> None:
    """model.py:102-106.  dense (B,D,H,W,C) f32, already zero."""
    _require_cuda(voxelwise, coords, dense)
    _contig(voxelwise, torch.float32), _contig(coords, torch.int64), _contig(dense, torch.float32)
    B, D, H, W, Cc = dense.shape
    with torch.cuda.device(dense.device):
        check(
            _cabi.lib().vn_scatter_dense_f32(
                _ptr(voxelwise), _ptr(coords), voxelwise.shape[0], _ptr(n_voxels_dev), Cc, B, D, H, W, _ptr(dense), _stream()
            ),
            "vn_scatter_dense_f32",
        )


def scatter_dense_bf16(voxelwise, coords, dense, clear: bool = False, n_voxels_dev=None) -> None:
    """Write (clear=False) or re-zero (clear=True) the occupied rows of the resident bf16 grid."""
    _require_cuda(voxelwise, coords, dense)
    _contig(coords, torch.int64), _contig(dense, torch.bfloat16)
    B, D, H, W, Cc = dense.shape
    with torch.cuda.device(dense.device):
        check(
            _cabi.lib().vn_scatter_dense_bf16(
                _ptr(voxelwise), _ptr(coords), coords.shape[0], _ptr(n_voxels_dev), Cc, B, D, H, W, _ptr(dense),
                1 if clear else 0, _stream(),
            ),
            "vn_scatter_dense_bf16",
        )


# ---------------------------------------------------------------------------------------------
# Conv3d middle
# ---------------------------------------------------------------------------------------------
def conv3d_pack_weights(weight, stride_d: int) -> torch.Tensor:
    """torch (64,cin,3,3,3) f32 -> packed bf16 tile image for a layer with depth stride `stride_d`
    (see vn_conv3d.cu)."""
    _require_cuda(weight)
    w = _contig(weight.detach(), torch.float32)
    cout, cin = w.shape[0], w.shape[1]
    if cout != 64 or tuple(w.shape[2:]) != (3, 3, 3):
        raise RuntimeError("conv3d kernel supports cout=64, 3x3x3 filters (model.py:207-209); got %s" % (tuple(w.shape),))
    nbytes = int(_cabi.lib().vn_conv3d_packed_weight_bytes(cin))
    if nbytes == 0:
        raise RuntimeError("conv3d kernel supports cin in {64,128}; got %d" % cin)
    packed = torch.empty(nbytes // 2, dtype=torch.bfloat16, device=w.device)
    with torch.cuda.device(w.device):
        check(_cabi.lib().vn_conv3d_pack_weights(_ptr(w), cin, stride_d, _ptr(packed), _stream()), "vn_conv3d_pack_weights")
    return packed


def conv3d_out_depth(din: int, stride_d: int, pad_d: int) -> int:
    return (din + 2 * pad_d - 3) // stride_d + 1


def conv3d_box_flags_bytes(B: int, Din: int, H: int, W: int) -> int:
    return int(_cabi.lib().vn_conv3d_box_flags_bytes(B, Din, H, W))


def conv3d_workspace_bytes(B: int, H: int, W: int) -> int:
    return int(_cabi.lib().vn_conv3d_workspace_bytes(B, H, W))


def conv3d_new_box_flags(B: int, Din: int, H: int, W: int, device) -> torch.Tensor:
    """Zeroed live-box table for a layer with input (B, Din, H, W, .)."""
    return torch.zeros(conv3d_box_flags_bytes(B, Din, H, W), dtype=torch.uint8, device=device)


def conv3d_new_out_state(B: int, H: int, W: int, device) -> torch.Tensor:
    """State table of a resident conv output buffer (see `out_state` in the header), all blocks stale."""
    n = int(_cabi.lib().vn_conv3d_out_state_bytes(B, H, W))
    return torch.full((n,), 0xFF, dtype=torch.uint8, device=device)


def conv3d_mark_boxes(coords, B: int, D: int, H: int, W: int, layers, flags, n_voxels_dev=None) -> None:
    """Flag the live input boxes of a chain of stacked middle layers from voxel coordinates (K,4) i64.
    layers: [(stride_d, pad_d), ...] (<= 3), flags: one uint8 table per layer (conv3d_new_box_flags)."""
    _require_cuda(coords, *flags)
    _contig(coords, torch.int64)
    n = len(layers)
    if n != len(flags) or not 1 <= n <= 3:
        raise RuntimeError("conv3d_mark_boxes: need 1..3 layers with one flag table each")
    din = D
    for (s, p), f in zip(layers, flags):
        if f.dtype != torch.uint8 or f.numel() != conv3d_box_flags_bytes(B, din, H, W):
            raise RuntimeError("conv3d_mark_boxes: flag table size does not match the layer input (B,%d,%d,%d)" % (din, H, W))
        din = conv3d_out_depth(din, s, p)
    strides = (C.c_int32 * n)(*[int(s) for s, _ in layers])
    pads = (C.c_int32 * n)(*[int(p) for _, p in layers])
    ptrs = (C.c_void_p * n)(*[f.data_ptr() for f in flags])
    with torch.cuda.device(coords.device):
        check(
            _cabi.lib().vn_conv3d_mark_boxes(_ptr(coords), coords.shape[0], _ptr(n_voxels_dev), B, D, H, W, n, strides, pads,
                                             ptrs, _stream()),
            "vn_conv3d_mark_boxes",
        )


def conv3d_fold_background(weight, din: int, stride_d: int, pad_d: int, scale, shift, bg_in=None, want_bg_out: bool = True):
    """Shift table (and output background) of a layer whose input is stored as a deviation from the background bg_in -
    see vn_conv3d_fold_background.  bg_in None (zero) or (Din,cin): -> (shift_table (3,3,Dout,64), bg_out (3,3,Dout,64));
    bg_in (3,3,Din,cin) (the bg_out of such a layer): -> (shift_table (5,5,Dout,64), None)."""
    _require_cuda(weight, scale, shift, bg_in)
    w = _contig(weight.detach(), torch.float32)
    cin = w.shape[1]
    dout = conv3d_out_depth(din, stride_d, pad_d)
    classes = 1
    if bg_in is not None:
        if bg_in.dtype != torch.float32 or not bg_in.is_contiguous() or tuple(bg_in.shape) not in ((din, cin), (3, 3, din, cin)):
            raise RuntimeError("bg_in must be a contiguous (Din, cin) or (3, 3, Din, cin) float32 tensor")
        classes = 9 if bg_in.dim() == 4 else 1
    nv = 5 if classes == 9 else 3
    table = torch.empty((nv, nv, dout, 64), dtype=torch.float32, device=w.device)
    bg_out = torch.empty((3, 3, dout, 64), dtype=torch.float32, device=w.device) if (want_bg_out and classes == 1) else None
    with torch.cuda.device(w.device):
        check(
            _cabi.lib().vn_conv3d_fold_background(_ptr(w), cin, din, stride_d, pad_d, _ptr(_contig(scale, torch.float32)),
                                                  _ptr(_contig(shift, torch.float32)), _ptr(bg_in), classes, _ptr(table), _ptr(bg_out),
                                                  _stream()),
            "vn_conv3d_fold_background",
        )
    return table, bg_out


def conv3d_plan(B: int, Din: int, H: int, W: int, stride_d: int, pad_d: int, box_flags, workspace, consume_flags: bool = False,
                border_ring: bool = False, out_state=None) -> None:
    """Build the work plan of a conv3d_bn_relu_fwd call ahead of time (it depends on the voxel coordinates only);
    pass the same box_flags / out_state / workspace and planned=True to the forward call."""
    _require_cuda(box_flags, workspace, out_state)
    opts = (_cabi.VN_CONV_FLAGS_CONSUME if consume_flags else 0) | (_cabi.VN_CONV_FLAGS_BORDER_RING if border_ring else 0)
    with torch.cuda.device(workspace.device):
        check(_cabi.lib().vn_conv3d_plan(B, Din, H, W, stride_d, pad_d, _ptr(box_flags), opts, _ptr(out_state), _ptr(workspace),
                                         workspace.numel(), _stream()), "vn_conv3d_plan")


def conv3d_bn_relu_fwd(x, packed_weight, scale, shift, stride_d: int, pad_d: int, final: bool = False, out=None,
                       box_flags=None, consume_flags: bool = False, border_ring: bool = False, sub=None, workspace=None,
                       out_state=None, planned: bool = False, reverse: bool = False):
    """x (B,Din,H,W,cin) bf16 NDHWC -> (B,Dout,H,W,64) bf16, or with final=True the reference's
    reshaped (B, 64*Dout, H, W) f32 (model.py:262), or with final="nhwc_bf16" a (B,H,W,64*Dout) bf16 tensor
    whose channel d*64 + c is the reference's channel c*Dout + d.

    shift: (64,) plain BatchNorm shift, (Dout,64), or the (3,3,Dout,64) table of conv3d_fold_background.
    sub: optional (Dout,64) background subtracted after the ReLU.  box_flags: optional live-box table
    (conv3d_mark_boxes); consume_flags resets it, border_ring forces boxes on the H/W border live.
    out_state: optional state table (conv3d_new_out_state) that makes `out` a resident buffer - only
    blocks fed now or by the previous call are written."""
    _require_cuda(x, packed_weight, scale, shift, box_flags, sub, workspace, out_state)
    if out_state is not None:
        if out is None:
            raise RuntimeError("out_state needs the resident `out` buffer it describes")
        if out_state.dtype != torch.uint8 or out_state.numel() != int(_cabi.lib().vn_conv3d_out_state_bytes(x.shape[0], x.shape[2], x.shape[3])):
            raise RuntimeError("out_state must be a uint8 tensor of vn_conv3d_out_state_bytes(B, H, W) elements")
    _contig(x, torch.bfloat16)
    B, Din, H, W, cin = x.shape
    if box_flags is not None and (box_flags.dtype != torch.uint8 or box_flags.numel() != conv3d_box_flags_bytes(B, Din, H, W)):
        raise RuntimeError("box_flags must be a uint8 tensor of conv3d_box_flags_bytes(B, Din, H, W) elements")
    Dout = conv3d_out_depth(Din, stride_d, pad_d)
    _contig(shift, torch.float32)
    if tuple(shift.shape) == (64,):
        n_variants = 0
    elif tuple(shift.shape) == (Dout, 64):
        n_variants = 1
    elif tuple(shift.shape) == (3, 3, Dout, 64):
        n_variants = 9
    elif tuple(shift.shape) == (5, 5, Dout, 64):
        n_variants = 25
    else:
        raise RuntimeError("shift must have shape (64,), (Dout,64), (3,3,Dout,64) or (5,5,Dout,64); got %s" % (tuple(shift.shape),))
    want_sub = (3, 3, Dout, 64) if n_variants == 9 else (Dout, 64)  # with a 9-variant table the background has 9 classes too
    if sub is not None and (tuple(sub.shape) != want_sub or sub.dtype != torch.float32 or not sub.is_contiguous()):
        raise RuntimeError("sub must be a contiguous %s float32 tensor (shift has %d variants); got %s" % (want_sub, n_variants, tuple(sub.shape)))
    if final not in (False, True, "nhwc_bf16"):
        raise RuntimeError("final must be False, True or 'nhwc_bf16'")
    out_mode = 2 if final == "nhwc_bf16" else (1 if final else 0)
    if out is None:
        if out_mode == 1:
            out = torch.empty((B, 64 * Dout, H, W), dtype=torch.float32, device=x.device)
        elif out_mode == 2:
            out = torch.empty((B, H, W, 64 * Dout), dtype=torch.bfloat16, device=x.device)
        else:
            out = torch.empty((B, Dout, H, W, 64), dtype=torch.bfloat16, device=x.device)
    want_out = [((B, Dout, H, W, 64), torch.bfloat16), ((B, 64 * Dout, H, W), torch.float32), ((B, H, W, 64 * Dout), torch.bfloat16)][out_mode]
    if tuple(out.shape) != want_out[0] or out.dtype != want_out[1] or not out.is_contiguous():
        raise RuntimeError("out must be a contiguous %s %s tensor; got %s %s" % (want_out[0], want_out[1], tuple(out.shape), out.dtype))
    if workspace is None:
        workspace = torch.empty(conv3d_workspace_bytes(B, H, W), dtype=torch.uint8, device=x.device)
    opts = (_cabi.VN_CONV_FLAGS_CONSUME if consume_flags else 0) | (_cabi.VN_CONV_FLAGS_BORDER_RING if border_ring else 0)
    if planned:
        opts |= _cabi.VN_CONV_FLAGS_PLANNED
    if reverse:  # walk the plan last group first (same bits): the blocks the previous layer wrote last are still in L2
        opts |= _cabi.VN_CONV_FLAGS_REVERSE
    with torch.cuda.device(x.device):
        check(
            _cabi.lib().vn_conv3d_bn_relu_fwd(
                _ptr(x), B, Din, H, W, cin, stride_d, pad_d, _ptr(packed_weight), _ptr(_contig(scale, torch.float32)),
                _ptr(shift), n_variants, _ptr(sub), out_mode, _ptr(out), _ptr(box_flags), opts,
                _ptr(out_state), _ptr(workspace), workspace.numel(), _stream(),
            ),
            "vn_conv3d_bn_relu_fwd",
        )
    return out


def conv3d_bn_relu_fwd_rows(grid, rows, voxel_offsets, n_points_total: int, vox_workspace, packed_weight, scale, shift,
                            stride_d: int, pad_d: int, out, box_flags=None, consume_flags: bool = False, sub=None,
                            workspace=None, out_state=None, planned: bool = False):
    """conv3d_bn_relu_fwd for the first middle layer (model.py:207) straight from the voxel-wise rows: rows (cap,128)
    bf16 from voxelize_svfe_rows, the occupancy index in `vox_workspace` (left by voxelize_index for the same grid,
    frame count and n_points_total) as the cell -> row map.  No (B,D,H,W,128) tensor exists; bit-identical to
    conv3d_bn_relu_fwd on the scattered grid.  out: (B,Dout,H,W,64) bf16."""
    _require_cuda(rows, voxel_offsets, vox_workspace, packed_weight, scale, shift, box_flags, sub, workspace, out_state, out)
    _contig(rows, torch.bfloat16), _contig(out, torch.bfloat16)
    B = voxel_offsets.numel() - 1
    Dout = conv3d_out_depth(grid.D, stride_d, pad_d)
    if tuple(out.shape) != (B, Dout, grid.H, grid.W, 64):
        raise RuntimeError("out must be (B,Dout,H,W,64) bf16; got %s" % (tuple(out.shape),))
    _contig(shift, torch.float32)
    if tuple(shift.shape) == (64,):
        n_variants = 0
    elif tuple(shift.shape) == (Dout, 64):
        n_variants = 1
    elif tuple(shift.shape) == (3, 3, Dout, 64):
        n_variants = 9
    else:
        raise RuntimeError("shift must have shape (64,), (Dout,64) or (3,3,Dout,64); got %s" % (tuple(shift.shape),))
    if sub is not None and (tuple(sub.shape) != (Dout, 64) or sub.dtype != torch.float32 or not sub.is_contiguous()):
        raise RuntimeError("sub must be a contiguous (Dout,64) float32 tensor")
    if box_flags is not None and (box_flags.dtype != torch.uint8 or box_flags.numel() != conv3d_box_flags_bytes(B, grid.D, grid.H, grid.W)):
        raise RuntimeError("box_flags must be a uint8 tensor of conv3d_box_flags_bytes(B, D, H, W) elements")
    if workspace is None:
        workspace = torch.empty(conv3d_workspace_bytes(B, grid.H, grid.W), dtype=torch.uint8, device=rows.device)
    opts = (_cabi.VN_CONV_FLAGS_CONSUME if consume_flags else 0) | (_cabi.VN_CONV_FLAGS_PLANNED if planned else 0)
    with torch.cuda.device(rows.device):
        check(
            _cabi.lib().vn_conv3d_bn_relu_fwd_rows(
                C.byref(grid), _ptr(rows), rows.shape[0], _ptr(voxel_offsets), B, int(n_points_total), _ptr(vox_workspace),
                vox_workspace.numel() * vox_workspace.element_size(), stride_d, pad_d, _ptr(packed_weight),
                _ptr(_contig(scale, torch.float32)), _ptr(shift), n_variants, _ptr(sub), 0, _ptr(out), _ptr(box_flags), opts,
                _ptr(out_state), _ptr(workspace), workspace.numel(), _stream(),
            ),
            "vn_conv3d_bn_relu_fwd_rows",
        )
    return out


# ---------------------------------------------------------------------------------------------
# 2-D 3x3 convolution (RPN blocks, model.py:212-251) on the same tcgen05 kernel
# ---------------------------------------------------------------------------------------------
def conv2d_workspace_bytes(B: int, H: int, W: int) -> int:
    return int(_cabi.lib().vn_conv2d_workspace_bytes(B, H, W))


def conv2d_pack_weights(weight, scale=None) -> torch.Tensor:
    """weight (cout,cin,3,3) f32 [* scale (cout,): BatchNorm folded in] -> the packed bf16 image of vn_conv2d_bn_relu_fwd."""
    _require_cuda(weight, scale)
    w = _contig(weight.detach(), torch.float32)
    cout, cin = int(w.shape[0]), int(w.shape[1])
    if tuple(w.shape[2:]) not in ((3, 3), (1, 1)):
        raise RuntimeError("conv2d_pack_weights: 3x3 or 1x1 kernels only; got %s" % (tuple(w.shape),))
    ksize = int(w.shape[2])
    nbytes = int(_cabi.lib().vn_conv2d_packed_weight_bytes(cin))
    if nbytes == 0:
        raise RuntimeError("conv2d_pack_weights: cin=%d unsupported (multiples of 64 up to 1024)" % cin)
    packed = torch.empty(nbytes, dtype=torch.uint8, device=w.device)
    with torch.cuda.device(w.device):
        check(_cabi.lib().vn_conv2d_pack_weights(_ptr(w), _ptr(None if scale is None else _contig(scale, torch.float32)), cin, cout,
                                                 ksize, _ptr(packed), _stream()), "vn_conv2d_pack_weights")
    return packed


def conv2d_bn_relu_fwd(x, packed_weight, shift, out=None, workspace=None, channel_offset: int = 0, planned: bool = False,
                       stride: int = 1) -> torch.Tensor:
    """x (B,H,W,cin) bf16 channels-last -> relu(conv3x3(x) + shift) as (B,Ho,Wo,cout) bf16 (padding 1; stride 1 or 2);
    shift (cout,) f32 carries the bias and the BatchNorm shift, the BatchNorm scale sits in the packed weights.
    `out` may have more channels than cout: the launch writes channels channel_offset .. + cout of it.  planned: `workspace`
    holds the plan of an earlier launch with the same B, Ho, Wo and cout - reuse it."""
    _require_cuda(x, packed_weight, shift, out, workspace)
    _contig(x, torch.bfloat16), _contig(shift, torch.float32)
    B, H, W, cin = x.shape
    Ho, Wo = ((H - 1) // 2 + 1, (W - 1) // 2 + 1) if stride == 2 else (H, W)
    cout = int(shift.numel())
    if out is None:
        out = torch.empty((B, Ho, Wo, cout), dtype=torch.bfloat16, device=x.device)
    elif tuple(out.shape[:3]) != (B, Ho, Wo) or out.dtype != torch.bfloat16 or not out.is_contiguous():
        raise RuntimeError("out must be a contiguous (B,Ho,Wo,C) bf16 tensor")
    if workspace is None:
        workspace = torch.empty(conv2d_workspace_bytes(B, Ho, Wo), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        check(_cabi.lib().vn_conv2d_bn_relu_fwd(_ptr(x), B, H, W, cin, cout, int(stride), _ptr(packed_weight), _ptr(shift), _ptr(out),
                                                int(out.shape[3]), int(channel_offset), 1 if planned else 0, _ptr(workspace),
                                                workspace.numel(), _stream()),
              "vn_conv2d_bn_relu_fwd")
    return out


def conv2d_pointwise_fwd(x, packed_weight, shift, out, channel_offset: int = 0, up: int = 1, oy: int = 0, ox: int = 0,
                         workspace=None, planned: bool = False, relu: bool = True) -> torch.Tensor:
    """1x1 convolution + shift + ReLU of x (B,H,W,cin) bf16 written to positions (h*up + oy, w*up + ox), channels
    channel_offset .. + cout, of out (B,H*up,W*up,C) bf16 - one of the up*up launches of a transposed convolution with
    kernel = stride = up (vn_conv2d_pointwise_fwd)."""
    _require_cuda(x, packed_weight, shift, out, workspace)
    _contig(x, torch.bfloat16), _contig(shift, torch.float32), _contig(out, torch.bfloat16)
    B, H, W, cin = x.shape
    if tuple(out.shape[:3]) != (B, H * up, W * up):
        raise RuntimeError("out must be (B, H*up, W*up, C); got %s" % (tuple(out.shape),))
    if workspace is None:
        workspace = torch.empty(conv2d_workspace_bytes(B, H, W), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        check(_cabi.lib().vn_conv2d_pointwise_fwd(_ptr(x), B, H, W, cin, int(shift.numel()), _ptr(packed_weight), _ptr(shift), 1 if relu else 0, _ptr(out),
                                                  int(out.shape[3]), int(channel_offset), int(up), int(oy), int(ox), 1 if planned else 0,
                                                  _ptr(workspace), workspace.numel(), _stream()), "vn_conv2d_pointwise_fwd")
    return out


def conv3d_plan_stats(workspace, B: int, Din: int, H: int, W: int, cin: int, stride_d: int, pad_d: int) -> dict:
    """Diagnostics (blocking, host side): decode the work plan the last vn_conv3d_bn_relu_fwd call left
    in `workspace` (PlanHeader 16 B | GroupDesc[n_tiles] 16 B each | TilePlan[n_tiles] 8 B each | per-group copies, see vn_conv3d.cu) and count
    what the kernel actually executed: live input boxes, tcgen05 FLOPs, operand bytes loaded by TMA.
    Used by bench.py for the executed-work roofline; not part of the product path."""
    import numpy as np

    _require_cuda(workspace)
    tiles_h, tiles_w = (H + 15) // 16, W // 8
    n_tiles = B * tiles_h * tiles_w
    raw = workspace.cpu().numpy()
    hdr = raw[:16].view(np.int32)
    tiles = raw[16 + 16 * n_tiles: 16 + 24 * n_tiles].view(np.uint32).reshape(n_tiles, 2)
    live = tiles[:, 0]
    touched = tiles[:, 1] & 0xFF
    restore = (tiles[:, 1] >> 24) & 0xFF
    Dout = conv3d_out_depth(Din, stride_d, pad_d)
    nkd = []
    for p in range(Din):
        n = 0
        for kd in range(3):
            num = p + pad_d - kd
            if num >= 0 and num % stride_d == 0 and num // stride_d < Dout:
                n += 1
        nkd.append(n)
    nchunk = cin // 64
    boxes = 0
    flops = 0
    live_planes = 0
    for p in range(Din):
        cnt = 0
        for kw in range(3):
            cnt += int(((live >> (p * 3 + kw)) & 1).sum())
        live_planes += int((((live >> (p * 3)) & 7) != 0).sum())
        boxes += cnt
        flops += cnt * nchunk * 12 * 2 * 128 * 64 * nkd[p] * 16
    popc = np.zeros(n_tiles, dtype=np.int64)
    rpop = np.zeros(n_tiles, dtype=np.int64)
    for d in range(8):
        popc += (touched >> d) & 1
        rpop += (restore >> d) & 1
    return {
        "tiles": int(n_tiles), "live_tiles": int((live != 0).sum()), "groups": int(hdr[0]),
        "live_boxes": int(boxes), "boxes_total": int(n_tiles * sum(1 for n in nkd if n) * 3),
        "touched_blocks": int(popc.sum()), "restored_blocks": int(rpop.sum()), "blocks_total": int(n_tiles * Dout),
        "executed_flops": int(flops), "a_bytes": int(boxes * nchunk * 18432),
        "live_tile_planes": live_planes, "unique_input_bytes": int(live_planes * 128 * cin * 2),
    }


# ---------------------------------------------------------------------------------------------
# training-mode BatchNorm (batch statistics)
# ---------------------------------------------------------------------------------------------
def vfe_layer_batch_stats(inputs, weight, bias) -> torch.Tensor:
    """Sums for the batch statistics of ONE stand-alone VFE layer (model.py:74-76 in train mode): inputs (K,T,cin) f32
    -> (2, units) f64, row 0 = sum over all K*T rows of relu(W x + b), row 1 = sum of squares."""
    _require_cuda(inputs, weight, bias)
    _contig(inputs, torch.float32)
    K, T, cin = inputs.shape
    units = weight.shape[0]
    sums = torch.empty((2, units), dtype=torch.float64, device=inputs.device)
    with torch.cuda.device(inputs.device):
        check(
            _cabi.lib().vn_vfe_layer_batch_stats(_ptr(inputs), K, T, cin, units, _ptr(_contig(weight, torch.float32)),
                                                 _ptr(_contig(bias, torch.float32)), _ptr(sums), _stream()),
            "vn_vfe_layer_batch_stats",
        )
    return sums


def svfe_batch_stats(feature, stage: int, w1, b1, s1=None, t1=None, w2=None, b2=None, n_voxels_dev=None) -> torch.Tensor:
    """Sums for the batch statistics of VFE-1 (stage 1 -> (2,16) f64) or VFE-2 (stage 2 -> (2,64) f64) over all
    K*T rows of feature (K,T,7): row 0 = sum of relu(Linear), row 1 = sum of squares (model.py:75-76)."""
    _require_cuda(feature, w1, b1, s1, t1, w2, b2)
    _contig(feature, torch.float32)
    K, T, cin = feature.shape
    if cin != 7:
        raise RuntimeError("svfe_batch_stats expects 7 input features per point")
    units = 16 if stage == 1 else 64
    sums = torch.zeros((2, units), dtype=torch.float64, device=feature.device)
    f32 = lambda t: None if t is None else _ptr(_contig(t, torch.float32))
    with torch.cuda.device(feature.device):
        check(
            _cabi.lib().vn_svfe_batch_stats(_ptr(feature), K, T, _ptr(n_voxels_dev), stage, f32(w1), f32(b1), f32(s1), f32(t1),
                                            f32(w2), f32(b2), _ptr(sums), _stream()),
            "vn_svfe_batch_stats",
        )
    return sums


def conv3d_raw_stats_fwd(x, packed_weight, bias, stride_d: int, pad_d: int, out=None, box_flags=None,
                         consume_flags: bool = False, workspace=None):
    """Training-mode ConvMD, first half (model.py:159): x (B,Din,H,W,cin) bf16 -> (raw (B,Dout,H,W,64) bf16 =
    conv(x)+bias, sums (2,64) f64 = per-channel sum / sum of squares of the fp32 raw output)."""
    _require_cuda(x, packed_weight, bias, box_flags, workspace)
    _contig(x, torch.bfloat16)
    B, Din, H, W, cin = x.shape
    Dout = conv3d_out_depth(Din, stride_d, pad_d)
    if out is None:
        out = torch.empty((B, Dout, H, W, 64), dtype=torch.bfloat16, device=x.device)
    if workspace is None:
        workspace = torch.empty(conv3d_workspace_bytes(B, H, W), dtype=torch.uint8, device=x.device)
    ones = torch.ones(64, dtype=torch.float32, device=x.device)
    sums = torch.zeros((2, 64), dtype=torch.float64, device=x.device)
    opts = _cabi.VN_CONV_FLAGS_CONSUME if consume_flags else 0
    with torch.cuda.device(x.device):
        check(
            _cabi.lib().vn_conv3d_raw_stats_fwd(_ptr(x), B, Din, H, W, cin, stride_d, pad_d, _ptr(packed_weight), _ptr(ones),
                                                _ptr(_contig(bias, torch.float32)), _ptr(out), _ptr(sums), _ptr(box_flags), opts,
                                                _ptr(workspace), workspace.numel(), _stream()),
            "vn_conv3d_raw_stats_fwd",
        )
    return out, sums


def bn_relu_apply(raw, scale, shift, final: bool = False, out=None):
    """Training-mode ConvMD, second half (model.py:161-165): relu(raw*scale+shift).  raw (B,D,H,W,64) bf16 ->
    bf16 same shape (in place when out is raw), or with final=True the (B, 64*D, H, W) f32 of model.py:262."""
    _require_cuda(raw, scale, shift, out)
    _contig(raw, torch.bfloat16)
    B, D, H, W, Cc = raw.shape
    if Cc != 64:
        raise RuntimeError("bn_relu_apply expects 64 channels")
    if out is None:
        out = torch.empty((B, 64 * D, H, W), dtype=torch.float32, device=raw.device) if final else torch.empty_like(raw)
    with torch.cuda.device(raw.device):
        check(
            _cabi.lib().vn_bn_relu_apply(_ptr(raw), B, D, H, W, _ptr(_contig(scale, torch.float32)),
                                         _ptr(_contig(shift, torch.float32)), 1 if final else 0, _ptr(out), _stream()),
            "vn_bn_relu_apply",
        )
    return out


# ---------------------------------------------------------------------------------------------
# backward of the Conv3d middle layers
# ---------------------------------------------------------------------------------------------
def conv3d_wgrad(x, dy, stride_d: int, pad_d: int, box_flags=None) -> torch.Tensor:
    """Weight gradient in torch layout (64, cin, 3, 3, 3) f32 from the layer input x (B,Din,H,W,cin) bf16 and the
    gradient dy (B,Dout,H,W,64) bf16 of the raw convolution output (tcgen05, fp32 accumulation).  box_flags: live-box
    table of x (conv3d_mark_boxes); dead boxes must be all-zero and are skipped."""
    _require_cuda(x, dy, box_flags)
    _contig(x, torch.bfloat16), _contig(dy, torch.bfloat16)
    B, Din, H, W, cin = x.shape
    Dout = conv3d_out_depth(Din, stride_d, pad_d)
    if tuple(dy.shape) != (B, Dout, H, W, 64):
        raise RuntimeError("dy must be (B,Dout,H,W,64) = %s, got %s" % ((B, Dout, H, W, 64), tuple(dy.shape)))
    taps = torch.zeros((27, cin, 64), dtype=torch.float32, device=x.device)
    with torch.cuda.device(x.device):
        check(_cabi.lib().vn_conv3d_wgrad(_ptr(x), _ptr(dy), B, Din, H, W, cin, stride_d, pad_d, _ptr(taps), _ptr(box_flags),
                                          _stream()),
              "vn_conv3d_wgrad")
    return taps.view(3, 3, 3, cin, 64).permute(4, 3, 0, 1, 2).contiguous()


def conv3d_pack_weights_dgrad(weight) -> torch.Tensor:
    """Packed bf16 image of the transposed, flipped filter for conv3d_dgrad (torch (64,64,3,3,3) f32 in)."""
    w = weight.detach()
    if w.shape[1] != 64:
        raise RuntimeError("conv3d_dgrad supports layers with 64 input channels (the first layer's data gradient is the sparse gather)")
    return conv3d_pack_weights(w.flip(2, 3, 4).transpose(0, 1).contiguous(), 1)


def conv3d_dgrad(dy, packed_weight_t, din_fwd: int, stride_d: int, pad_d: int, workspace=None) -> torch.Tensor:
    """dx (B,Din_fwd,H,W,64) bf16 from dy (B,Dout_fwd,H,W,64) bf16 (gradient of the raw convolution output)."""
    _require_cuda(dy, packed_weight_t, workspace)
    _contig(dy, torch.bfloat16)
    B, Dout, H, W, Cc = dy.shape
    if Cc != 64 or conv3d_out_depth(din_fwd, stride_d, pad_d) != Dout:
        raise RuntimeError("dy shape %s does not match a layer with Din=%d stride_d=%d pad_d=%d" % (tuple(dy.shape), din_fwd, stride_d, pad_d))
    dx = torch.empty((B, din_fwd, H, W, 64), dtype=torch.bfloat16, device=dy.device)
    if workspace is None:
        workspace = torch.empty(conv3d_workspace_bytes(B, H, W), dtype=torch.uint8, device=dy.device)
    ones = torch.ones(64, dtype=torch.float32, device=dy.device)
    zeros = torch.zeros(64, dtype=torch.float32, device=dy.device)
    scratch = torch.zeros(128, dtype=torch.float64, device=dy.device)
    with torch.cuda.device(dy.device):
        check(
            _cabi.lib().vn_conv3d_dgrad(_ptr(dy), B, Dout, H, W, din_fwd, stride_d, pad_d, _ptr(packed_weight_t), _ptr(ones),
                                        _ptr(zeros), _ptr(dx), _ptr(scratch), _ptr(workspace), workspace.numel(), _stream()),
            "vn_conv3d_dgrad",
        )
    return dx


# ---------------------------------------------------------------------------------------------
# backward of BatchNorm(batch statistics) + ReLU, of to_dense + conv-1 (sparse), and of the stacked VFE
# ---------------------------------------------------------------------------------------------
def bn_relu_bwd(raw, dy, scale, shift, mean, invstd, final: bool = False):
    """Backward of relu(BatchNorm3d(raw)) in training mode (model.py:161-165).  raw (B,D,H,W,64) bf16; dy is the
    gradient of the layer output: bf16 (B,D,H,W,64), or with final=True f32 (B,64*D,H,W) (model.py:262 layout).
    Returns (draw (B,D,H,W,64) bf16 = gradient of the raw convolution output, dgamma (64,) f32, dbeta (64,) f32)."""
    _require_cuda(raw, dy, scale, shift, mean, invstd)
    _contig(raw, torch.bfloat16)
    _contig(dy, torch.float32 if final else torch.bfloat16)
    B, D, H, W, Cc = raw.shape
    if Cc != 64:
        raise RuntimeError("bn_relu_bwd expects 64 channels")
    want = (B, 64 * D, H, W) if final else (B, D, H, W, 64)
    if tuple(dy.shape) != want:
        raise RuntimeError("dy must be %s, got %s" % (want, tuple(dy.shape)))
    n = B * D * H * W
    prm = torch.zeros((6, 64), dtype=torch.float32, device=raw.device)
    prm[0], prm[1], prm[2], prm[3] = scale, shift, mean, invstd
    sums = torch.zeros((2, 64), dtype=torch.float64, device=raw.device)
    draw = torch.empty_like(raw)
    mode = 1 if final else 0
    with torch.cuda.device(raw.device):
        check(_cabi.lib().vn_bn_relu_bwd(_ptr(raw), _ptr(dy), mode, B, D, H, W, _ptr(prm), 0, _ptr(sums), None, _stream()),
              "vn_bn_relu_bwd(reduce)")
        prm[4:6] = (sums / n).float()
        check(_cabi.lib().vn_bn_relu_bwd(_ptr(raw), _ptr(dy), mode, B, D, H, W, _ptr(prm), 1, None, _ptr(draw), _stream()),
              "vn_bn_relu_bwd(apply)")
    return draw, sums[1].float(), sums[0].float()


def conv1_weight_taps(weight) -> torch.Tensor:
    """(64,128,3,3,3) f32 -> [27][128 c][64 n] bf16: the values the bf16 forward multiplied with, one K-major
    B tile per tap for the gather-GEMM."""
    w = weight.detach().to(torch.bfloat16)
    return w.permute(2, 3, 4, 1, 0).reshape(27, w.shape[1], w.shape[0]).contiguous()


def conv1_dgrad_gather(draw, weight_taps, coords, stride_d: int, pad_d: int) -> torch.Tensor:
    """Gradient of the (K,128) voxel-wise features through to_dense + the first conv layer: draw (B,Dr,H,W,64) bf16,
    weight_taps from conv1_weight_taps, coords (K,4) i64 -> (K,128) f32."""
    _require_cuda(draw, weight_taps, coords)
    _contig(draw, torch.bfloat16), _contig(weight_taps, torch.bfloat16), _contig(coords, torch.int64)
    B, Dr, H, W, Cc = draw.shape
    if Cc != 64 or tuple(weight_taps.shape) != (27, 128, 64):
        raise RuntimeError("conv1_dgrad_gather: draw (B,Dr,H,W,64) and weight_taps (27,128,64) expected")
    K = coords.shape[0]
    dvw = torch.empty((K, 128), dtype=torch.float32, device=draw.device)
    with torch.cuda.device(draw.device):
        check(_cabi.lib().vn_conv1_dgrad_gather(_ptr(draw), _ptr(weight_taps), _ptr(coords), K, B, Dr, H, W, stride_d, pad_d,
                                                _ptr(dvw), _stream()), "vn_conv1_dgrad_gather")
    return dvw


def svfe_bwd(feature, dvw, w1, b1, w2, b2, bn1, bn2) -> dict:
    """Backward of the stacked VFE in training mode.  feature (K,T,7) f32, dvw (K,128) f32; bn1 / bn2 =
    (scale, shift, mean, invstd) of the two batch-statistics BatchNorms as used by the forward.
    Returns {'w1','b1','gamma1','beta1','w2','b2','gamma2','beta2'} gradients (f32)."""
    _require_cuda(feature, dvw, w1, b1, w2, b2, *bn1, *bn2)
    _contig(feature, torch.float32), _contig(dvw, torch.float32)
    K, T, cin = feature.shape
    if cin != 7 or tuple(dvw.shape) != (K, 128):
        raise RuntimeError("svfe_bwd: feature (K,T,7) and dvw (K,128) expected")
    dev = feature.device
    p1 = torch.zeros((6, 16), dtype=torch.float32, device=dev)
    p2 = torch.zeros((6, 64), dtype=torch.float32, device=dev)
    for i in range(4):
        p1[i], p2[i] = bn1[i], bn2[i]
    acc = torch.zeros(2400, dtype=torch.float64, device=dev)
    rows = torch.empty((K, T, 16), dtype=torch.float32, device=dev)  # only the evaluated rows are touched
    n = float(max(K * T, 1))
    args = [_ptr(_contig(t, torch.float32)) for t in (w1, b1, w2, b2)]
    with torch.cuda.device(dev):
        def run(p):
            check(_cabi.lib().vn_svfe_bwd(_ptr(feature), K, T, _ptr(dvw), p, *args, _ptr(p1), _ptr(p2), _ptr(rows), _ptr(acc), _stream()),
                  "vn_svfe_bwd(pass %d)" % p)
        run(1)
        p2[4], p2[5] = (acc[2176:2240] / n).float(), (acc[2112:2176] / n).float()
        run(2)
        p1[4], p1[5] = (acc[2384:2400] / n).float(), (acc[2368:2384] / n).float()
        run(3)
    f = acc.float()
    return {
        "w2": f[0:2048].view(64, 32), "b2": f[2048:2112], "gamma2": f[2112:2176], "beta2": f[2176:2240],
        "w1": f[2240:2352].view(16, 7), "b1": f[2352:2368], "gamma1": f[2368:2384], "beta1": f[2384:2400],
    }


# ---------------------------------------------------------------------------------------------
# RPN target assignment
# ---------------------------------------------------------------------------------------------
def bbox_iou(box1, box2) -> torch.Tensor:
    """(N,4) f32 x (K,4) f32 stand-up boxes -> (N,K) f32 overlaps, bit-identical to utils.py:344-373."""
    _require_cuda(box1, box2)
    _contig(box1, torch.float32), _contig(box2, torch.float32)
    if box1.dim() != 2 or box2.dim() != 2 or box1.shape[1] != 4 or box2.shape[1] != 4:
        raise RuntimeError("bbox_iou expects (N,4) and (K,4) boxes; got %s, %s" % (tuple(box1.shape), tuple(box2.shape)))
    out = torch.zeros((box1.shape[0], box2.shape[0]), dtype=torch.float32, device=box1.device)
    with torch.cuda.device(box1.device):
        check(
            _cabi.lib().vn_bbox_iou(_ptr(box1), _ptr(box2), box1.shape[0], box2.shape[0], _ptr(out), _stream()),
            "vn_bbox_iou",
        )
    return out


def rpn_targets(anchors_standup, anchors, gt_standup, gt_boxes, gt_offsets, pos_iou: float, neg_iou: float,
                anchor_h: float):
    """One batch of generate_targets (see header).  Returns (pos, neg) (B,A) f64 and targets (B,A,7) f64."""
    _require_cuda(anchors_standup, anchors, gt_standup, gt_boxes, gt_offsets)
    _contig(anchors_standup, torch.float32), _contig(anchors, torch.float64)
    _contig(gt_standup, torch.float32), _contig(gt_boxes, torch.float64), _contig(gt_offsets, torch.int32)
    A, B, G = anchors.shape[0], gt_offsets.numel() - 1, gt_boxes.shape[0]
    if anchors_standup.shape != (A, 4) or anchors.shape != (A, 7) or gt_standup.shape != (G, 4) or gt_boxes.shape != (G, 7):
        raise RuntimeError("rpn_targets: inconsistent shapes")
    dev = anchors.device
    pos = torch.empty((B, A), dtype=torch.float64, device=dev)
    neg = torch.empty((B, A), dtype=torch.float64, device=dev)
    targets = torch.empty((B, A, 7), dtype=torch.float64, device=dev)
    ws = torch.empty(int(_cabi.lib().vn_rpn_targets_workspace_bytes(B, A, G)), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        check(
            _cabi.lib().vn_rpn_targets(
                _ptr(anchors_standup), _ptr(anchors), A, _ptr(gt_standup) if G else None, _ptr(gt_boxes) if G else None,
                _ptr(gt_offsets), B, G, pos_iou, neg_iou, anchor_h, _ptr(pos), _ptr(neg), _ptr(targets), _ptr(ws),
                ws.numel(), _stream(),
            ),
            "vn_rpn_targets",
        )
    return pos, neg, targets


def rpn_targets_static_capacity_bytes() -> int:
    return int(_cabi.lib().vn_rpn_targets_static_capacity_bytes())


def rpn_targets_static(anchors_standup, anchors, gt_standup, gt_boxes, gt_offsets, pos_iou: float, neg_iou: float,
                       anchor_h: float, pos, neg, targets):
    """`rpn_targets` through vn_rpn_targets_static: the scratch is the library's static device array and the CALLER passes
    the outputs (pos, neg (B,A) f64; targets (B,A,7) f64), so nothing is allocated on this path - neither by the library
    nor here.  Raises when the batch needs more scratch than the array holds (header: VN_RPN_TARGETS_STATIC_BYTES).
    Calls on one device share the array: do not let two of them overlap on the GPU."""
    _require_cuda(anchors_standup, anchors, gt_standup, gt_boxes, gt_offsets, pos, neg, targets)
    _contig(anchors_standup, torch.float32), _contig(anchors, torch.float64)
    _contig(gt_standup, torch.float32), _contig(gt_boxes, torch.float64), _contig(gt_offsets, torch.int32)
    _contig(pos, torch.float64), _contig(neg, torch.float64), _contig(targets, torch.float64)
    A, B, G = anchors.shape[0], gt_offsets.numel() - 1, gt_boxes.shape[0]
    if anchors_standup.shape != (A, 4) or anchors.shape != (A, 7) or gt_standup.shape != (G, 4) or gt_boxes.shape != (G, 7):
        raise RuntimeError("rpn_targets_static: inconsistent shapes")
    if pos.numel() != B * A or neg.numel() != B * A or targets.numel() != B * A * 7:
        raise RuntimeError("rpn_targets_static: outputs must hold (B,A), (B,A) and (B,A,7) float64")
    with torch.cuda.device(anchors.device):
        check(
            _cabi.lib().vn_rpn_targets_static(
                _ptr(anchors_standup), _ptr(anchors), A, _ptr(gt_standup) if G else None, _ptr(gt_boxes) if G else None,
                _ptr(gt_offsets), B, G, pos_iou, neg_iou, anchor_h, _ptr(pos), _ptr(neg), _ptr(targets), _stream(),
            ),
            "vn_rpn_targets_static",
        )
    return pos, neg, targets
