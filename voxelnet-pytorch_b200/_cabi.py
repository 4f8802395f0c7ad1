"""ctypes binding of include/voxelnet_b200.h.

`lib()` loads `lib/libvoxelnet_b200.so` (building it first if it is missing and nvcc is around)
and declares every prototype the header declares.  There is no fallback: if the library cannot
be loaded the import error is raised to the caller, and every compute entry point needs a CUDA
device.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libvoxelnet_b200.so")

VN_ABI_VERSION = 16
VN_CONV_FLAGS_CONSUME = 1
VN_CONV_FLAGS_BORDER_RING = 2
VN_CONV_FLAGS_PLANNED = 4
VN_CONV_FLAGS_REVERSE = 8


class VnGrid(C.Structure):
    """struct vn_grid (voxelnet/utils.py:24-33 constants)."""

    _fields_ = [
        ("D", C.c_int32), ("H", C.c_int32), ("W", C.c_int32), ("T", C.c_int32),
        ("off_x", C.c_float), ("off_y", C.c_float), ("off_z", C.c_float),
        ("vs_x", C.c_float), ("vs_y", C.c_float), ("vs_z", C.c_float),
    ]


_p = C.c_void_p
_i32, _i64, _sz = C.c_int32, C.c_int64, C.c_size_t
_G = C.POINTER(VnGrid)

# name -> (restype, argtypes): must list every symbol of include/voxelnet_b200.h
PROTOTYPES = {
    "vn_abi_version": (C.c_int, []),
    "vn_last_error": (C.c_char_p, []),
    "vn_grid_for_target": (C.c_int, [C.c_char_p, _G]),
    "vn_voxelize_workspace_bytes": (_sz, [_G, _i32, _i64]),
    "vn_voxelize_index": (C.c_int, [_G, _p, _p, _i32, _i64, _i32, _p, _p, _sz, _p]),
    "vn_voxelize_fill": (C.c_int, [_G, _p, _p, _i32, _i64, _i32, _p, _i64, _p, _p, _p, _p, _sz, _p]),
    "vn_voxelize_counts_host": (C.c_int, [_p, _i32, _p, _p]),
    "vn_vfe_layer_fwd": (C.c_int, [_p, _p, _i64, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p]),
    "vn_vfe_layer_batch_stats": (C.c_int, [_p, _i64, _i32, _i32, _i32, _p, _p, _p, _p]),
    "vn_feature_mask": (C.c_int, [_p, _i64, _i32, _i32, _p, _p]),
    "vn_svfe_fused_fwd": (C.c_int, [_p, _i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "vn_voxelize_svfe_fused": (C.c_int, [_G, _p, _p, _i32, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vn_voxelize_svfe_bucketed": (C.c_int, [_G, _p, _p, _i32, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vn_voxelize_svfe_rows": (C.c_int, [_G, _p, _p, _i32, _i64, _p, _i64, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _sz, _p]),
    "vn_conv3d_bn_relu_fwd_rows": (C.c_int, [_G, _p, _i64, _p, _i32, _i64, _p, _sz, _i32, _i32, _p, _p, _p, _i32, _p, _i32, _p, _p, _i32, _p, _p,
                                             _sz, _p]),
    "vn_voxelize_bucket_coords": (C.c_int, [_G, _p, _p, _i32, _i64, _p, _i64, _p, _p, _sz, _p]),
    "vn_fill_zero": (C.c_int, [_p, _sz, _p]),
    "vn_scatter_dense_f32": (C.c_int, [_p, _p, _i64, _p, _i32, _i32, _i32, _i32, _i32, _p, _p]),
    "vn_scatter_dense_bf16": (C.c_int, [_p, _p, _i64, _p, _i32, _i32, _i32, _i32, _i32, _p, _i32, _p]),
    "vn_conv3d_packed_weight_bytes": (_sz, [_i32]),
    "vn_conv3d_pack_weights": (C.c_int, [_p, _i32, _i32, _p, _p]),
    "vn_conv3d_box_flags_bytes": (_sz, [_i32, _i32, _i32, _i32]),
    "vn_conv3d_workspace_bytes": (_sz, [_i32, _i32, _i32]),
    "vn_conv3d_out_state_bytes": (_sz, [_i32, _i32, _i32]),
    "vn_conv3d_bn_relu_fwd": (C.c_int, [_p, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _p, _p, _p, _i32, _p, _i32, _p, _p, _i32, _p, _p, _sz, _p]),
    "vn_bbox_iou": (C.c_int, [_p, _p, _i64, _i32, _p, _p]),
    "vn_rpn_targets_workspace_bytes": (_sz, [_i32, _i32, _i32]),
    "vn_rpn_targets": (C.c_int, [_p, _p, _i32, _p, _p, _p, _i32, _i32, C.c_float, C.c_float, C.c_double, _p, _p, _p,
                                 _p, _sz, _p]),
    "vn_rpn_targets_static_capacity_bytes": (_sz, []),
    "vn_rpn_targets_static": (C.c_int, [_p, _p, _i32, _p, _p, _p, _i32, _i32, C.c_float, C.c_float, C.c_double, _p, _p, _p, _p]),
    "vn_conv2d_workspace_bytes": (_sz, [_i32, _i32, _i32]),
    "vn_conv2d_packed_weight_bytes": (_sz, [_i32]),
    "vn_conv2d_pack_weights": (C.c_int, [_p, _p, _i32, _i32, _i32, _p, _p]),
    "vn_conv2d_pointwise_fwd": (C.c_int, [_p, _i32, _i32, _i32, _i32, _i32, _p, _p, _i32, _p, _i32, _i32, _i32, _i32, _i32, _i32, _p, _sz, _p]),
    "vn_conv2d_bn_relu_fwd": (C.c_int, [_p, _i32, _i32, _i32, _i32, _i32, _i32, _p, _p, _p, _i32, _i32, _i32, _p, _sz, _p]),
    "vn_conv3d_plan": (C.c_int, [_i32, _i32, _i32, _i32, _i32, _i32, _p, _i32, _p, _p, _sz, _p]),
    "vn_conv3d_fold_background": (C.c_int, [_p, _i32, _i32, _i32, _i32, _p, _p, _p, _i32, _p, _p, _p]),
    "vn_conv3d_mark_boxes": (C.c_int, [_p, _i64, _p, _i32, _i32, _i32, _i32, _i32, _p, _p, _p, _p]),
    "vn_conv3d_raw_stats_fwd": (C.c_int, [_p, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p, _i32, _p, _sz, _p]),
    "vn_bn_relu_apply": (C.c_int, [_p, _i32, _i32, _i32, _i32, _p, _p, _i32, _p, _p]),
    "vn_svfe_batch_stats": (C.c_int, [_p, _i64, _i32, _p, _i32, _p, _p, _p, _p, _p, _p, _p, _p]),
    "vn_conv3d_dgrad": (C.c_int, [_p, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _p, _p, _p, _p, _p, _p, _sz, _p]),
    "vn_conv3d_wgrad": (C.c_int, [_p, _p, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _p, _p, _p]),
    "vn_bn_relu_bwd": (C.c_int, [_p, _p, _i32, _i32, _i32, _i32, _i32, _p, _i32, _p, _p, _p]),
    "vn_conv1_dgrad_gather": (C.c_int, [_p, _p, _p, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _p, _p]),
    "vn_svfe_bwd": (C.c_int, [_p, _i64, _i32, _p, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    "vn_launch_count": (_i64, []),
    "vn_host_pack_threads": (C.c_int, [_i32]),
    "vn_host_pack_submit": (_i64, [_p, _p, _i32, _p, _i64, _p]),
    "vn_host_pack_wait": (C.c_int, [_i64]),
    "vn_upload_frames": (C.c_int, [_p, _p, _i32, _p, _i64, _p, _p, _p]),
}

_lib = None
_lock = threading.Lock()


class VnError(RuntimeError):
    """A libvoxelnet_b200 entry point returned non-zero."""


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            from . import build as _build  # builds in-tree; raises if nvcc is missing

            _build.build()
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(handle, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        got = handle.vn_abi_version()
        if got != VN_ABI_VERSION:
            raise ImportError("libvoxelnet_b200.so ABI %d != expected %d (stale build?)" % (got, VN_ABI_VERSION))
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().vn_last_error()
        raise VnError("%s failed (rc=%d): %s" % (what, rc, msg.decode() if msg else "?"))


def grid_for_target(target: str) -> VnGrid:
    g = VnGrid()
    check(lib().vn_grid_for_target(target.encode(), C.byref(g)), "vn_grid_for_target")
    return g


def launch_count() -> int:
    return int(lib().vn_launch_count())
