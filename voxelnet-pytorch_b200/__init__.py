"""voxelnet_pytorch_b200 - the VoxelNet front-end hot path (voxelize -> stacked VFE -> sparse-to-dense
-> Conv3d middle layers) as hand-written sm_100a CUDA kernels behind the reference's Python
interfaces.  See DESIGN.md for the design and INTEGRATION.md for how it drops into
johanngerberding/voxelnet-pytorch.

Importing the package does not load the CUDA library; the first call into `ops` does, and fails
loudly if `lib/libvoxelnet_b200.so` is missing and cannot be built, or if no CUDA device exists.
"""
from .config import CAR, PED_CYC, GridConfig, grid_config  # noqa: F401
from .dataset import PinnedFrameBatch, collate_fn, load_velodyne_bin, pinned_frame, prepare_voxel  # noqa: F401
from .model import ConvMD, FeatureLearningNet, MiddleConvNet, VFELayer  # noqa: F401
from .pipeline import HostFrameStream, VoxelNetFrontEnd  # noqa: F401
from .rpn import ConvMD2d, DeConvMD2d, DeConvUpMD2d, HybridRPN, MiddleRPN, RPN, RPNBlockTail  # noqa: F401
from .targets import TargetAssigner, bbox_iou, generate_anchors, generate_targets, label_to_gt_box_3d  # noqa: F401
from .utils import VoxelBatch, pack_frames, pcl_to_voxels, voxelize_batch  # noqa: F401

__all__ = [
    "CAR", "PED_CYC", "GridConfig", "grid_config", "collate_fn", "prepare_voxel", "ConvMD", "FeatureLearningNet",
    "MiddleConvNet", "VFELayer", "ConvMD2d", "DeConvMD2d", "DeConvUpMD2d", "HybridRPN", "MiddleRPN", "RPN", "RPNBlockTail", "VoxelNetFrontEnd", "HostFrameStream", "PinnedFrameBatch", "load_velodyne_bin", "pinned_frame", "VoxelBatch", "pack_frames", "pcl_to_voxels", "voxelize_batch",
    "TargetAssigner", "bbox_iou", "generate_anchors", "generate_targets", "label_to_gt_box_3d",
]
