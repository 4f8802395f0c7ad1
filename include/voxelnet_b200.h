/*
 * voxelnet_b200.h - C ABI of libvoxelnet_b200.so: the VoxelNet front-end hot path
 * (voxelize -> stacked VFE -> sparse-to-dense scatter -> Conv3d middle layers) as hand-written
 * sm_100a CUDA kernels.
 *
 * The reference (johanngerberding/voxelnet-pytorch) is pure Python and has no FFI; the interfaces
 * these entry points replace are Python call sites, cited per function as `file:line` relative to
 * the reference root.  The Python host side that mirrors those call sites lives in
 * voxelnet-pytorch_b200/{utils,dataset,model}.py and binds this header with ctypes
 * (voxelnet-pytorch_b200/_cabi.py); INTEGRATION.md shows the stub a reference maintainer adds.
 *
 * Conventions
 *   - Every pointer is a DEVICE pointer unless its name ends in `_host`.  The caller allocates all
 *     buffers including workspaces; the library owns no device memory.
 *   - `stream` is a cudaStream_t passed as void*.  Calls only enqueue work: no hidden host
 *     synchronisation, no host callbacks.  (The one exception is documented at vn_voxelize_counts_host.)
 *   - Return value: 0 = ok, < 0 = bad argument (VN_ERR_*), > 0 = cudaError_t.  vn_last_error()
 *     returns a thread-local message for the last non-zero return.
 *   - There is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef VOXELNET_B200_H
#define VOXELNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VN_ABI_VERSION 16

#define VN_OK 0
#define VN_ERR_BAD_ARG (-1)
#define VN_ERR_WORKSPACE (-2)
#define VN_ERR_UNSUPPORTED (-3)
#define VN_ERR_DRIVER (-4)

typedef void* vn_stream_t; /* cudaStream_t */

/* Per-class grid constants: voxelnet/utils.py:24-33 (voxelizer) and voxelnet/config.py:38-50,61-73
 * (model side).  Axis naming follows the reference: D = z cells, H = y cells, W = x cells. */
typedef struct vn_grid {
  int32_t D, H, W;       /* grid_size (z, y, x): Car 10x400x352, Ped/Cyc 10x200x240          */
  int32_t T;             /* max points per voxel: 35 / 45                                     */
  float off_x, off_y, off_z; /* lidar_coord added to (x, y, z): (0, 40, 3) / (0, 20, 3)       */
  float vs_x, vs_y, vs_z;    /* voxel size along (x, y, z): (0.2, 0.2, 0.4), as float32       */
} vn_grid_t;

int vn_abi_version(void);
const char* vn_last_error(void);

/* utils.py:24-33: target == "Car" selects the Car grid, any other string Pedestrian/Cyclist. */
int vn_grid_for_target(const char* target, vn_grid_t* grid_out);

/* ------------------------------------------------------------------------------------------
 * Voxelizer - replaces pcl_to_voxels (voxelnet/utils.py:10-100) and prepare_voxel's batch column
 * (voxelnet/dataset.py:101-119) for a whole batch of frames in one call sequence.
 *
 * Frames are concatenated: points[frame_offsets[f] .. frame_offsets[f+1]) belong to frame f,
 * each point 4 x float32 (x, y, z, reflectance) as in dataset.py:51.  Point ORDER is the
 * caller's (the reference's shuffle, utils.py:35, is the caller's business - see DESIGN.md).
 *
 * Phase 1 (vn_voxelize_index): range crop + voxel index with IEEE fp32 add/divide/floor
 *   (utils.py:37-61), per-frame occupancy bitmap, sorted-unique rank of every occupied cell
 *   (== np.unique(axis=0) order, utils.py:63), per-voxel point counts and per-point arrival
 *   slots.  Writes voxel_offsets[0..n_frames] = exclusive prefix of voxels per frame
 *   (voxel_offsets[n_frames] = total K).
 * Phase 2 (vn_voxelize_fill): first-T-points-in-array-order selection (utils.py:78-84), 7-d
 *   features with fp64 centroid arithmetic (utils.py:87-88), int64 coordinates [b, z, y, x],
 *   int64 counts.  Rows >= max_voxels are dropped (the caller sized the outputs from
 *   voxel_offsets, or passed a capacity).
 * ------------------------------------------------------------------------------------------ */
size_t vn_voxelize_workspace_bytes(const vn_grid_t* grid, int32_t n_frames, int64_t n_points_total);

int vn_voxelize_index(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets,
                      int32_t n_frames, int64_t n_points_total, int32_t max_points_per_frame,
                      int32_t* voxel_offsets /* [n_frames + 1] */, void* workspace,
                      size_t workspace_bytes, vn_stream_t stream);

int vn_voxelize_fill(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets,
                     int32_t n_frames, int64_t n_points_total, int32_t max_points_per_frame,
                     const int32_t* voxel_offsets, int64_t max_voxels,
                     float* feature_buffer /* [max_voxels, T, 7]  utils.py:71     */,
                     int64_t* coordinate_buffer /* [max_voxels, 4] b,z,y,x  dataset.py:110-117 */,
                     int64_t* number_buffer /* [max_voxels]        utils.py:69     */,
                     void* workspace, size_t workspace_bytes, vn_stream_t stream);

/* Convenience for the NumPy-returning wrapper: synchronises `stream` and copies
 * voxel_offsets[0..n_frames] to the host.  This is the only entry point that blocks. */
int vn_voxelize_counts_host(const int32_t* voxel_offsets, int32_t n_frames,
                            int32_t* voxel_offsets_host, vn_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Stacked VFE - replaces VFELayer.forward (voxelnet/model.py:74-82) and the VFE part of
 * FeatureLearningNet.forward (model.py:95-100).  Eval-mode BatchNorm1d is passed folded as a
 * post-ReLU affine: y = relu(W x + b) * bn_scale + bn_shift (Linear -> ReLU -> BN, model.py:68-76).
 * ------------------------------------------------------------------------------------------ */

/* One VFE layer, drop-in shapes: inputs [K, T, cin] f32, mask [K, T] u8 (bool) ->
 * out [K, T, 2*units] f32 = concat(pointwise, max over all T rows) * mask.
 * Supported (cin, units): (7, 16) and (32, 64) - the two instances at model.py:88-89. */
int vn_vfe_layer_fwd(const float* inputs, const uint8_t* mask, int64_t K, int32_t T, int32_t cin,
                     int32_t units, const float* weight /* [units, cin] */,
                     const float* bias /* [units] */, const float* bn_scale, const float* bn_shift,
                     float* out, vn_stream_t stream);

/* Training-mode statistics of ONE stand-alone layer (VFELayer.forward under .train(), model.py:74-76): the
 * BatchNorm1d there normalises relu(Linear(x)) with the statistics of all K*T rows (pad rows included).
 * sums [2][units] f64 <- per-unit sum and sum of squares of relu(W x + b); zeroed by the call.  The caller folds
 * them into (bn_scale, bn_shift) for vn_vfe_layer_fwd and updates the running statistics. */
int vn_vfe_layer_batch_stats(const float* inputs, int64_t K, int32_t T, int32_t cin, int32_t units,
                             const float* weight, const float* bias, double* sums /* [2][units] */,
                             vn_stream_t stream);

/* model.py:95-96: mask[k, t] = (max_j feature[k, t, j] != 0). */
int vn_feature_mask(const float* feature, int64_t K, int32_t T, int32_t cin, uint8_t* mask,
                    vn_stream_t stream);

/* Fused VFE-1 -> VFE-2 -> voxel-wise max (model.py:95-100): feature [K, T, 7] f32 ->
 * voxelwise [K, 128] f32.  The mask is derived from the features exactly as the reference does.
 * `n_voxels_dev` (optional, may be NULL) points at a device int32 holding the live K (<= K);
 * it lets a pipeline run without a host read of K. */
int vn_svfe_fused_fwd(const float* feature, int64_t K, int32_t T, const int32_t* n_voxels_dev,
                      const float* w1 /* [16,7] */, const float* b1, const float* s1, const float* t1,
                      const float* w2 /* [64,32] */, const float* b2, const float* s2,
                      const float* t2, float* voxelwise /* [K,128] */, vn_stream_t stream);

/* Fully fused fast path: phase 2 of the voxelizer and the stacked VFE in one kernel.  After
 * vn_voxelize_index, builds each voxel's distinct feature rows on chip (first-T selection, fp64
 * centroid - utils.py:78-88) and evaluates VFE-1 -> VFE-2 -> voxel-wise max on them (model.py:95-100).
 * Outputs: coordinate_buffer [max_voxels,4] i64 (number_buffer may be NULL) and the voxel-wise maxima,
 * either as voxelwise [max_voxels,128] f32, or written straight into their cells of the resident bf16
 * grid dense_bf16 [n_frames, D, H, W, 128] (the scatter of model.py:102-106, same rounding as
 * vn_scatter_dense_bf16), or both; one of the two may be NULL.  The [K,T,7] feature buffer is never
 * materialised; results are bit-identical to vn_voxelize_fill + vn_svfe_fused_fwd (+ scatter). */
int vn_voxelize_svfe_fused(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets,
                           int32_t n_frames, int64_t n_points_total, const int32_t* voxel_offsets,
                           int64_t max_voxels, const float* w1, const float* b1, const float* s1,
                           const float* t1, const float* w2, const float* b2, const float* s2,
                           const float* t2, float* voxelwise, void* dense_bf16,
                           int64_t* coordinate_buffer, int64_t* number_buffer, void* workspace,
                           size_t workspace_bytes, vn_stream_t stream);

/* The same in two calls, for pipelines that want the voxel coordinates early: vn_voxelize_bucket_coords does the
 * bucketing and writes coordinate_buffer [max_voxels,4] (everything that depends on the coordinates only - live-box
 * marking, conv planning - can then run on another stream next to the SVFE kernel); vn_voxelize_svfe_bucketed is
 * vn_voxelize_svfe_fused without the bucketing step and must follow it on the same workspace; it takes the same
 * coordinate_buffer but does not write it again (another stream may be reading it). */
int vn_voxelize_bucket_coords(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets,
                              int32_t n_frames, int64_t n_points_total, const int32_t* voxel_offsets,
                              int64_t max_voxels, int64_t* coordinate_buffer, void* workspace,
                              size_t workspace_bytes, vn_stream_t stream);
int vn_voxelize_svfe_bucketed(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets,
                           int32_t n_frames, int64_t n_points_total, const int32_t* voxel_offsets,
                           int64_t max_voxels, const float* w1, const float* b1, const float* s1,
                           const float* t1, const float* w2, const float* b2, const float* s2,
                           const float* t2, float* voxelwise, void* dense_bf16,
                           int64_t* coordinate_buffer, int64_t* number_buffer, void* workspace,
                           size_t workspace_bytes, vn_stream_t stream);

/* The never-materialised form of the sparse -> dense step (voxelnet/model.py:102-106, consumer :207): the 128 voxel-wise
 * maxima of voxel row r go to rows_bf16[r][0..127] (bf16, RNE) - a compact (max_voxels,128) table in np.unique row
 * order - and NO (B,D,H,W,128) tensor is written.  The cell -> row map is the voxelizer's own occupancy index inside
 * `workspace` (bitmap + popcount prefix, see vn_conv3d_bn_relu_fwd_rows), so nothing else is stored.  bucketed != 0: the
 * call follows vn_voxelize_bucket_coords on the same workspace (coordinate_buffer is then not written again); 0: it
 * does the bucketing itself and writes coordinate_buffer. */
int vn_voxelize_svfe_rows(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets,
                          int32_t n_frames, int64_t n_points_total, const int32_t* voxel_offsets,
                          int64_t max_voxels, const float* w1, const float* b1, const float* s1,
                          const float* t1, const float* w2, const float* b2, const float* s2,
                          const float* t2, void* rows_bf16, int64_t* coordinate_buffer, int32_t bucketed,
                          void* workspace, size_t workspace_bytes, vn_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Sparse -> dense - replaces torch.sparse.FloatTensor(coords.t(), voxelwise, [B,D,H,W,128])
 * .to_dense() (voxelnet/model.py:102-106).  Layout of `dense`: [B, D, H, W, C] contiguous.
 * ------------------------------------------------------------------------------------------ */
int vn_fill_zero(void* ptr, size_t bytes, vn_stream_t stream);

/* fp32 drop-in: dense must be zero-filled by the caller; duplicates are summed (COO coalesce). */
int vn_scatter_dense_f32(const float* voxelwise, const int64_t* coords /* [K,4] b,z,y,x */, int64_t K,
                         const int32_t* n_voxels_dev, int32_t C, int32_t B, int32_t D, int32_t H,
                         int32_t W, float* dense, vn_stream_t stream);

/* bf16 fast path: writes (or, with clear != 0, re-zeroes) the K occupied rows of a bf16 grid that
 * is kept resident and zero between frames, so the 360 MB/frame zero-fill is paid once. */
int vn_scatter_dense_bf16(const float* voxelwise, const int64_t* coords, int64_t K,
                          const int32_t* n_voxels_dev, int32_t C, int32_t B, int32_t D, int32_t H,
                          int32_t W, void* dense_bf16, int32_t clear, vn_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Conv3d middle layers - replaces ConvMD(3, ...).forward = Conv3d(k=3) -> BatchNorm3d -> ReLU
 * (voxelnet/model.py:158-167; the three instances at :207-209) and the reshape at :262.
 * bf16 implicit GEMM on tcgen05/TMEM, TMA-fed; fp32 accumulate; eval-mode BN + conv bias folded
 * into a per-channel fp32 (scale, shift) applied in the epilogue before ReLU.
 * ------------------------------------------------------------------------------------------ */

/* Repack torch-layout fp32 weights [64, cin, 3, 3, 3] into the kernel's bf16 tap-major,
 * 128B-swizzled tile layout.  packed holds 27 * cin * 64 bf16.  The image depends on the depth
 * stride of the layer it will be used with (depth taps that share an input plane sit side by side). */
size_t vn_conv3d_packed_weight_bytes(int32_t cin);
int vn_conv3d_pack_weights(const float* weight, int32_t cin, int32_t stride_d, void* packed,
                           vn_stream_t stream);

/* Sizes of the two caller-allocated side buffers of the convolution:
 *   box flags  [B][Din][ceil(H/16)][W/8][3] bytes, one per 18x8 halo'd input box (see below); must be
 *              zero before the first vn_conv3d_mark_boxes;
 *   workspace  the per-call work plan (tile masks, tile groups, work counter); contents are scratch. */
size_t vn_conv3d_box_flags_bytes(int32_t B, int32_t Din, int32_t H, int32_t W);
size_t vn_conv3d_workspace_bytes(int32_t B, int32_t H, int32_t W);
/*   out state  [B][ceil(H/16)][W/8] bytes kept next to an OUTPUT buffer that is reused from call to call
 *              (see `out_state` below); fill with 0xFF when the buffer is new or the parameters changed. */
size_t vn_conv3d_out_state_bytes(int32_t B, int32_t H, int32_t W);

#define VN_CONV_FLAGS_CONSUME 1     /* reset every flag that was read, leaving the table all-zero again */
#define VN_CONV_FLAGS_BORDER_RING 2 /* also treat boxes that contain an H/W border position as live    */
#define VN_CONV_FLAGS_PLANNED 4     /* the work plan in `workspace` was built by vn_conv3d_plan already    */
#define VN_CONV_FLAGS_REVERSE 8     /* walk the plan last group first (same result; use it on every other layer of a
                                       chain: the blocks the previous layer wrote last are still in L2)          */

/* x: [B, Din, H, W, cin] bf16 (NDHWC).  cin in {64, 128}, cout = 64, kernel 3x3x3, H/W stride 1 and
 * padding 1; depth stride/padding (stride_d, pad_d) in {(2,1), (1,0)}; Din <= 10, W % 8 == 0.
 *
 *   y = relu(conv(x) * scale[n] + shift) - sub
 *
 * out_mode 0: y [B, Dout, H, W, 64] bf16 (NDHWC) - feeds the next layer.
 * out_mode 1: y [B, 64*Dout, H, W] f32 with channel = c*Dout + d - the reshape of model.py:262.
 * out_mode 2: y [B, H, W, 64*Dout] bf16 (channels last) with channel = d*64 + c - for a bf16 channels-last consumer
 *             (permute the input channels of its first convolution once: reference channel c*Dout + d <- d*64 + c).
 * shift_table / n_variants: n_variants 0: a plain [64] vector (BatchNorm shift with the conv bias
 *   folded in) - the drop-in case.  n_variants 1: [Dout][64].  n_variants 9: [3][3][Dout][64] from
 *   vn_conv3d_fold_background, indexed by (h class, w class) of the output position
 *   (0 interior, 1 first row/column, 2 last row/column).  n_variants 25: [5][5][Dout][64] (see
 *   vn_conv3d_fold_background with bg_classes 9).
 * sub (optional): [Dout][64] - or [3][3][Dout][64] when n_variants is 9 - subtracted after the ReLU: the layer's own
 *   background, so that the output is stored as a deviation that is exactly zero wherever the input was all background.
 * box_flags (optional): flag[b][p][th][tw][kw] == 0 asserts that the input box
 *   x[b, p, 16*th-1 .. 16*th+16, 8*tw+kw-1 .. 8*tw+kw+6, :] is all zeros; its load and its MMAs are
 *   skipped (exact: zero operands contribute nothing).  NULL: every box is live.  flag_opts: VN_CONV_FLAGS_*.
 * out_state (optional): makes y a RESIDENT buffer.  Where no live box reaches a 16x8 output block the
 *   layer writes a constant that depends only on its parameters (0 for a deviation-coded layer).  Bit d
 *   of out_state[b][th][tw] records that block (b, d, th, tw) of y currently holds anything else.  The
 *   call then writes only the blocks that are fed now (setting the bit) or were fed by the previous
 *   call on the same y (rewriting the constant, clearing the bit) - exact, and the untouched ~60 % of
 *   a KITTI-shaped frame is neither computed nor stored.  Start from all 0xFF (every block is
 *   written once).  NULL: every block of y is written.
 * workspace: vn_conv3d_workspace_bytes(B, H, W) bytes, 16-byte aligned. */
int vn_conv3d_bn_relu_fwd(const void* x, int32_t B, int32_t Din, int32_t H, int32_t W, int32_t cin,
                          int32_t stride_d, int32_t pad_d, const void* packed_weight,
                          const float* scale /* [64] */, const float* shift_table, int32_t n_variants,
                          const float* sub, int32_t out_mode, void* y, uint8_t* box_flags,
                          int32_t flag_opts, uint8_t* out_state, void* workspace,
                          size_t workspace_bytes, vn_stream_t stream);

/* vn_conv3d_bn_relu_fwd for the FIRST middle layer (voxelnet/model.py:207) reading the voxel-wise rows directly:
 * x is never a tensor.  Input cell (b, z, y, x) is rows_bf16[r] if bit (z*H + y)*W + x of frame b's occupancy bitmap is
 * set, r = voxel_offsets[b] + (set bits before it in the frame) < min(voxel_offsets[n_frames], max_voxels), else zero.
 * The bitmap and its per-word popcount prefix are the ones vn_voxelize_index left in vox_workspace for
 * (grid, n_frames, n_points_total) - pass the same three values; rows_bf16 comes from vn_voxelize_svfe_rows.  The
 * kernel's box-filling warps build each 18 x 8 x 64-channel operand box in shared memory from the few occupied rows
 * (everything else stays zero), so neither the 360 MB/frame grid nor its scatter / re-zero passes exist.  B = n_frames,
 * Din = grid->D, cin = 128; every other argument as in vn_conv3d_bn_relu_fwd.  Results are bit-identical to running
 * vn_conv3d_bn_relu_fwd on the scattered grid. */
int vn_conv3d_bn_relu_fwd_rows(const vn_grid_t* grid, const void* rows_bf16, int64_t max_voxels,
                               const int32_t* voxel_offsets, int32_t n_frames, int64_t n_points_total,
                               const void* vox_workspace, size_t vox_workspace_bytes, int32_t stride_d,
                               int32_t pad_d, const void* packed_weight, const float* scale,
                               const float* shift_table, int32_t n_variants, const float* sub,
                               int32_t out_mode, void* y, uint8_t* box_flags, int32_t flag_opts,
                               uint8_t* out_state, void* workspace, size_t workspace_bytes,
                               vn_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * 2-D 3x3 convolution + BatchNorm2d(eval) + ReLU on the same tcgen05 kernel - the stride-1 ConvMD(2, 128, 128, 3, (1,1),
 * (1,1)) layers of the RPN blocks (voxelnet/model.py:212-251).  Channels-last bf16 in and out:
 *   x [B, H, W, cin] -> y [B, H, W, cout],  y = relu(conv(x, weight * scale) + shift),  cin a multiple of 64 (<= 1024), cout in
 *   {64, 128, 192} per launch, any W (the last 8-wide tile column may be ragged).  stride 1, or 2 (the heads of the
 *   blocks, model.py:213,233,244): y is then [B, (H-1)/2+1, (W-1)/2+1, cout]; workspace and plan are those of the OUTPUT extent.  The groups of 64 output channels ride along the MMA's N dimension the way the depth
 *   taps of the 3-D layers do (one 128 x cout x 16 MMA per k-step).
 * vn_conv2d_pack_weights: weight [cout][cin][3][3] f32 (torch layout), scale [cout] = BatchNorm gamma / sqrt(var + eps) or
 *   NULL; packed: vn_conv2d_packed_weight_bytes(cin) bytes.  shift [cout] = beta + (bias - mean) * scale.
 * y has y_channels channels per position and the launch writes channels y_channel_offset .. + cout of it (a 256-channel
 *   layer is two launches of 128 into one tensor; ConvTranspose2d(k 3, stride 1, padding 1) = this convolution with the
 *   kernel flipped and its channel axes swapped: voxelnet/model.py:229 deconv1).
 * planned != 0: `workspace` still holds the work plan of an earlier launch with the same B, H, W and cout (the plan of a dense
 *   2-D layer depends on nothing else) - it is reused instead of rebuilt.
 * workspace: vn_conv2d_workspace_bytes(B, output H, output W).  The strided deconvolutions (kernel = stride) and the 1x1 heads go through
 *   vn_conv2d_pointwise_fwd below. */
size_t vn_conv2d_workspace_bytes(int32_t B, int32_t H, int32_t W);
size_t vn_conv2d_packed_weight_bytes(int32_t cin);
int vn_conv2d_pack_weights(const float* weight, const float* scale, int32_t cin, int32_t cout,
                           int32_t ksize /* 3, or 1 for vn_conv2d_pointwise_fwd */, void* packed,
                           vn_stream_t stream);
int vn_conv2d_bn_relu_fwd(const void* x, int32_t B, int32_t H, int32_t W, int32_t cin, int32_t cout,
                          int32_t stride, const void* packed_weight, const float* shift, void* y, int32_t y_channels,
                          int32_t y_channel_offset, int32_t planned, void* workspace,
                          size_t workspace_bytes, vn_stream_t stream);

/* 1x1 convolution (+ folded BatchNorm + ReLU) whose outputs are placed on an `up` times finer grid: position (h, w) of
 * x [B, H, W, cin] is written to (h * up + oy, w * up + ox) of y [B, H * up, W * up, y_channels], channels y_channel_offset ..
 * + cout.  A transposed convolution with kernel = stride = up and no padding (the reference's deconv2 / deconv3,
 * voxelnet/model.py:242,253: DeConv2d(128, 256, 2, (2,2), (0,0)) and DeConv2d(256, 256, 4, (4,4), (0,0))) is up * up such launches
 * per block of <= 192 output channels, launch (oy, ox) carrying the weights W[:, :, oy, ox] (packed with ksize 1: weight
 * [cout][cin] f32) - each output is written once, straight into its slice of the concatenated tensor, BatchNorm and ReLU in
 * the epilogue.  up = 1 is a plain 1x1 convolution; relu = 0 drops the activation (the RPN's prob_conv / reg_conv heads,
 * model.py:253-254: 768 -> 2 + 14 channels, padded to 64 output channels by the caller).  cin: a multiple of 64 up to 1024.  workspace / planned as in vn_conv2d_bn_relu_fwd (plans of 3x3 and 1x1
 * launches differ: do not share a planned workspace between them). */
int vn_conv2d_pointwise_fwd(const void* x, int32_t B, int32_t H, int32_t W, int32_t cin, int32_t cout,
                            const void* packed_weight, const float* shift, int32_t relu, void* y,
                            int32_t y_channels, int32_t y_channel_offset, int32_t up, int32_t oy, int32_t ox, int32_t planned,
                            void* workspace, size_t workspace_bytes, vn_stream_t stream);

/* Background folding for a layer whose INPUT is stored as a deviation from a background (what the previous
 * layer outputs wherever ITS input was all background; NULL = zero background).  The reference zero-pads, so
 * out-of-image taps see 0 rather than the background: the constant the background contributes to an output depends on
 * the output depth and on the border class of the output position.
 *   bg_classes 1: bg_in [Din][cin], the same at every position.  Writes shift_table [3][3][Dout][64] (= bias*scale +
 *     shift, for n_variants 9; classes: 0 interior, 1 first, 2 last row / column) and, if bg_out is not NULL, the
 *     layer's own output background per class bg_out [3][3][Dout][64] = relu(shift_table) - pass it as `sub`
 *     (n_variants 9) so that the stored deviation is exactly zero wherever no voxel is in reach, borders included.
 *   bg_classes 9: bg_in [3][3][Din][cin] is such a per-class background.  Writes shift_table [5][5][Dout][64] (for
 *     n_variants 25; classes 3 / 4: second / last but one row or column, which still see the border row of the input);
 *     bg_out must be NULL.
 * weight is the torch-layout fp32 tensor; it is rounded to bf16 exactly as the packed image is. */
int vn_conv3d_fold_background(const float* weight, int32_t cin, int32_t Din, int32_t stride_d,
                              int32_t pad_d, const float* scale, const float* shift,
                              const float* bg_in, int32_t bg_classes, float* shift_table, float* bg_out,
                              vn_stream_t stream);

/* Live-box flags for a chain of n_layers (<= 3) stacked middle layers from the voxel coordinates
 * [K,4] (b,z,y,x) of the scattered grid: a voxel can disturb the input of layer l only at rows
 * y-l..y+l, columns x-l..x+l and the planes its depth taps reach; every box containing such a
 * position is flagged in flags[l] ([B][Din_l][ceil(H/16)][W/8][3], Din_0 = D).  `flags` is a HOST
 * array of n_layers device pointers; stride_d / pad_d are host arrays of the layers' depth geometry. */
int vn_conv3d_mark_boxes(const int64_t* coords, int64_t K, const int32_t* n_voxels_dev, int32_t B,
                         int32_t D, int32_t H, int32_t W, int32_t n_layers,
                         const int32_t* stride_d_host, const int32_t* pad_d_host,
                         uint8_t* const* flags_host, vn_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Training-mode BatchNorm (model.train(True), voxelnet/train.py:148): batch statistics.
 * ------------------------------------------------------------------------------------------ */

/* The raw convolution for a ConvMD in training mode (model.py:159): y_raw = conv(x) + bias as bf16
 * NDHWC [B, Dout, H, W, 64], no BatchNorm, no ReLU; and, from the fp32 accumulators, the per-channel
 * sums BatchNorm3d's batch statistics need: stats[0..63] += sum(y), stats[64..127] += sum(y^2) over all
 * B*Dout*H*W positions (fp64, zeroed by the caller).  `ones` is a [64] vector of 1.0f.  box_flags /
 * flag_opts as in vn_conv3d_bn_relu_fwd (only exact for a zero-background input: the first layer). */
int vn_conv3d_raw_stats_fwd(const void* x, int32_t B, int32_t Din, int32_t H, int32_t W, int32_t cin,
                            int32_t stride_d, int32_t pad_d, const void* packed_weight,
                            const float* ones, const float* bias, void* y_raw, double* stats,
                            uint8_t* box_flags, int32_t flag_opts, void* workspace,
                            size_t workspace_bytes, vn_stream_t stream);

/* BatchNorm3d + ReLU applied to a stored raw convolution output (model.py:161-165) with the folded
 * per-channel scale/shift: out = relu(raw * scale[c] + shift[c]).  raw: bf16 [B, D, H, W, 64].
 * out_mode 0: bf16 [B, D, H, W, 64] (may alias raw); out_mode 1: f32 [B, 64*D, H, W], channel = c*D + d
 * (the reshape of model.py:262). */
int vn_bn_relu_apply(const void* raw_bf16, int32_t B, int32_t D, int32_t H, int32_t W, const float* scale,
                     const float* shift, int32_t out_mode, void* out, vn_stream_t stream);

/* Batch statistics of the two VFE BatchNorm1d layers (model.py:72,76) over all K*T rows of feature
 * [K, T, 7], pad rows included.  stage 1: sums[0..15] += sum relu(W1 x + b1), sums[16..31] += sum of
 * squares.  stage 2 (s1, t1 = VFE-1's BN folded from its batch statistics): sums[0..63] / sums[64..127]
 * likewise for relu(W2 [masked concat] + b2).  sums: fp64, zeroed by the caller. */
int vn_svfe_batch_stats(const float* feature, int64_t K, int32_t T, const int32_t* n_voxels_dev,
                        int32_t stage, const float* w1, const float* b1, const float* s1, const float* t1,
                        const float* w2, const float* b2, double* sums, vn_stream_t stream);

/* The planning step of vn_conv3d_bn_relu_fwd on its own: reads (and with VN_CONV_FLAGS_CONSUME resets) the live-box
 * table, updates the output state table and writes the work plan into `workspace`.  It depends on the voxel
 * coordinates only, not on the layer input, so the plans of all three layers can be built while an earlier layer
 * still computes (another stream / a parallel branch of a CUDA graph); the forward call then gets the same
 * box_flags / out_state / workspace and flag_opts | VN_CONV_FLAGS_PLANNED. */
int vn_conv3d_plan(int32_t B, int32_t Din, int32_t H, int32_t W, int32_t stride_d, int32_t pad_d,
                   uint8_t* box_flags, int32_t flag_opts, uint8_t* out_state, void* workspace,
                   size_t workspace_bytes, vn_stream_t stream);

/* Data gradient of a middle layer with 64 input channels (layers 2 and 3): dx [B, Din_fwd, H, W, 64] bf16 from
 * dy [B, Dout_fwd, H, W, 64] bf16 - the forward tcgen05 kernel run on the transposed depth geometry.
 * packed_weight_t = vn_conv3d_pack_weights(W.flip(kd,kh,kw).transpose(0,1), cin 64, stride_d 1).  ones / zeros:
 * [64] f32 vectors; stats_scratch: [128] f64 scratch (contents undefined afterwards). */
int vn_conv3d_dgrad(const void* dy, int32_t B, int32_t Dout_fwd, int32_t H, int32_t W, int32_t Din_fwd,
                    int32_t stride_d, int32_t pad_d, const void* packed_weight_t, const float* ones,
                    const float* zeros, void* dx, double* stats_scratch, void* workspace,
                    size_t workspace_bytes, vn_stream_t stream);

/* Weight gradient of a middle layer (autograd of nn.Conv3d, model.py:145-151 / train.py:149), bf16 operands on
 * tcgen05 with fp32 accumulation:  x [B, Din, H, W, cin] bf16 (the layer input), dy [B, Dout, H, W, 64] bf16 (the
 * gradient of the raw convolution output).  dw_taps [27][cin][64] f32, tap = (kd*3 + kh)*3 + kw, is ACCUMULATED
 * into (zero it first):  dw_taps[tap][c][n] += sum_pos dy[pos][n] * x[pos + tap][c].
 * box_flags (optional): the live-box table of x as written by vn_conv3d_mark_boxes; boxes flagged dead must hold
 * only zeros (true for the voxel grid of the first layer) and are skipped - exact, ~10x less work there. */
int vn_conv3d_wgrad(const void* x, const void* dy, int32_t B, int32_t Din, int32_t H, int32_t W, int32_t cin,
                    int32_t stride_d, int32_t pad_d, float* dw_taps, const uint8_t* box_flags,
                    vn_stream_t stream);

/* BatchNorm3d(batch statistics) + ReLU backward of a ConvMD layer (autograd of model.py:161-165, driven by
 * loss.backward() at train.py:149) over the stored raw convolution output raw [B, D, H, W, 64] bf16.
 * dy: dy_mode 0 = bf16 [B, D, H, W, 64]; dy_mode 1 = f32 [B, 64*D, H, W], channel = c*D + d (gradient of the
 * tensor reshaped at model.py:262).  prm: [6][64] f32 = scale (gamma*invstd), shift, mean, invstd, dbeta/N,
 * dgamma/N.  apply == 0: sums[0..63] += sum dz, sums[64..127] += sum dz*xhat (fp64; = dbeta, dgamma), with
 * dz = dy * (raw*scale + shift > 0); rows 4, 5 of prm are not read.  apply != 0: draw [B, D, H, W, 64] bf16 =
 * scale * (dz - prm[4] - xhat*prm[5]), the gradient of the raw convolution output. */
int vn_bn_relu_bwd(const void* raw_bf16, const void* dy, int32_t dy_mode, int32_t B, int32_t D, int32_t H,
                   int32_t W, const float* prm, int32_t apply, double* sums, void* draw_bf16,
                   vn_stream_t stream);

/* Backward of to_dense (model.py:102-106) fused with the data gradient of the first middle layer: the
 * transposed convolution is evaluated only at the K occupied cells.  draw [B, Dr, H, W, 64] bf16 = gradient of
 * that layer's raw output; weight_taps [27][128][64] bf16 = W[n][c][kd][kh][kw] as [tap][c][n]; coords [K,4]
 * (b, z, y, x); dvw [K,128] f32 = gradient of the voxel-wise features.  A gather-GEMM on tcgen05: 128 voxels x
 * 128 channels per CTA, K = 64 per valid tap, fp32 accumulation in TMEM. */
int vn_conv1_dgrad_gather(const void* draw_bf16, const void* weight_taps, const int64_t* coords, int64_t K,
                          int32_t B, int32_t Dr, int32_t H, int32_t W, int32_t stride_d, int32_t pad_d,
                          float* dvw, vn_stream_t stream);

/* Backward of the stacked VFE with batch-statistics BatchNorm1d (autograd of model.py:74-82, 95-100) from the
 * gradient dvw [K,128] of the voxel-wise features; feature [K, T, 7] as in vn_svfe_fused_fwd.  Three passes
 * (each recomputes the forward on chip): pass 1 -> dgamma2, dbeta2; pass 2 (needs bn2 rows 4, 5) -> dW2, db2,
 * dgamma1, dbeta1, and the VFE-1 output gradient of every evaluated row into row_scratch [K][T][16] f32 (caller-
 * allocated, may be NULL for pass 1); pass 3 (needs bn1 rows 4, 5 and row_scratch) -> dW1, db1.  bn1 [6][16], bn2 [6][64] f32 = scale, shift,
 * mean, invstd, dbeta/N, dgamma/N with N = K*T.  acc: 2400 fp64, accumulated into: dW2[64][32] @0, db2 @2048,
 * dgamma2 @2112, dbeta2 @2176, dW1[16][7] @2240, db1 @2352, dgamma1 @2368, dbeta1 @2384. */
int vn_svfe_bwd(const float* feature, int64_t K, int32_t T, const float* dvw, int32_t pass, const float* w1,
                const float* b1, const float* w2, const float* b2, const float* bn1, const float* bn2,
                float* row_scratch, double* acc, vn_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * RPN target assignment - replaces the interpreter loops of bbox_iou
 * (voxelnet/utils.py:344-373) and the per-frame body of generate_targets (utils.py:395-462).
 * ------------------------------------------------------------------------------------------ */
/* overlaps [N,K] f32 of box1 [N,4] and box2 [K,4] (x1, y1, x2, y2; f32, 16-byte aligned), float32 arithmetic in
 * the reference's operation order, bit-identical to utils.bbox_iou (including its anchor-area term as written at
 * utils.py:366-367).  N == 0 or K == 0 is a no-op. */
int vn_bbox_iou(const float* box1, const float* box2, int64_t N, int32_t K, float* overlaps, vn_stream_t stream);

size_t vn_rpn_targets_workspace_bytes(int32_t n_frames, int32_t n_anchors, int32_t n_boxes_total);

/* One batch of generate_targets.  anchors_standup [A,4] f32 and gt_standup [G,4] f32 are the float32 stand-up boxes
 * the reference hands to bbox_iou (utils.py:397-409); anchors [A,7] f64 and gt_boxes [G,7] f64 are (x,y,z,h,w,l,r);
 * gt_offsets [n_frames+1] i32 is the CSR of the boxes per frame (a frame may have none: all its anchors are
 * negative).  pos_iou / neg_iou = cfg.OBJECT.RPN_POS_IOU / RPN_NEG_IOU as float32 (the dtype numpy compares in),
 * anchor_h = cfg.OBJECT.ANCHOR_H.  Outputs, every element written: pos_equal_one, neg_equal_one [n_frames, A] f64
 * (= [B, H, W, 2]) and targets [n_frames, A, 7] f64 (= [B, H, W, 14]).  Index sets are identical to the
 * reference's; target values are IEEE-exact except the three log() terms (1 ulp). */
int vn_rpn_targets(const float* anchors_standup, const double* anchors, int32_t n_anchors, const float* gt_standup,
                   const double* gt_boxes, const int32_t* gt_offsets, int32_t n_frames, int32_t n_boxes_total,
                   float pos_iou, float neg_iou, double anchor_h, double* pos_equal_one, double* neg_equal_one,
                   double* targets, void* workspace, size_t workspace_bytes, vn_stream_t stream);

/* vn_rpn_targets without a caller workspace and without any allocation.  This entry point has no counterpart in the
 * reference: it was added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  The scratch of the four kernels lives in a
 * file-scope `static __device__` array of VN_RPN_TARGETS_STATIC_BYTES (16 MiB) inside the library; every GPU that loads
 * the library holds it until the process exits.  Same arguments, outputs and results as vn_rpn_targets (bit-identical);
 * the call enqueues the same four launches on `stream` and calls no cudaMalloc / cudaFree or allocating routine.
 * Limit: vn_rpn_targets_workspace_bytes(n_frames, n_anchors, n_boxes_total) <= VN_RPN_TARGETS_STATIC_BYTES, i.e.
 * about n_frames * n_anchors * 8 + n_boxes_total * 8 bytes - Car (70,400 anchors) up to 29 frames per call; beyond it
 * the call returns VN_ERR_WORKSPACE with the sizes in vn_last_error() and launches nothing.
 * Every call on a device shares the one array, which keeps the previous call's contents: the part that must start from
 * a known value is re-initialised by the call's first kernel, but two calls that OVERLAP on the GPU (different streams,
 * no ordering between them) race - serialise them, or use vn_rpn_targets with separate workspaces.
 * vn_rpn_targets_static_capacity_bytes() returns VN_RPN_TARGETS_STATIC_BYTES as compiled into the library. */
#define VN_RPN_TARGETS_STATIC_BYTES (16u << 20)
size_t vn_rpn_targets_static_capacity_bytes(void);
int vn_rpn_targets_static(const float* anchors_standup, const double* anchors, int32_t n_anchors, const float* gt_standup,
                          const double* gt_boxes, const int32_t* gt_offsets, int32_t n_frames, int32_t n_boxes_total,
                          float pos_iou, float neg_iou, double anchor_h, double* pos_equal_one, double* neg_equal_one,
                          double* targets, vn_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Host-side frame packer (HOST pointers only; no CUDA call) - the head of the input path that replaces the
 * reference's per-frame host voxelization + per-element `.to(device)` (voxelnet/dataset.py:51-58,
 * voxelnet/model.py:302-303): a list of per-frame (N_i, 4) float32 arrays is gathered into ONE (page-locked) point
 * buffer plus its CSR frame offsets by a pool of worker threads, so that a single asynchronous copy ships the batch.
 *
 * vn_host_pack_threads(n): sets the pool size before its first use (n > 0; default 4 or $VN_HOST_THREADS) and
 *   returns the size in effect (0 = not created yet and no size requested).
 * vn_host_pack_submit: enqueues the copy and returns a ticket > 0 at once (frame_offsets_host is filled before it
 *   returns; the frame arrays and the destination must stay valid until the ticket is waited for), or a VN_ERR_*.
 * vn_host_pack_wait(ticket): blocks until that job and every job submitted before it are complete.
 * ------------------------------------------------------------------------------------------ */
int vn_host_pack_threads(int32_t n_threads);
int64_t vn_host_pack_submit(const float* const* frames_host, const int64_t* n_points_host, int32_t n_frames,
                            float* points_host, int64_t capacity_points, int32_t* frame_offsets_host /* [n_frames+1] */);
int vn_host_pack_wait(int64_t ticket);

/* The same batch WITHOUT the host-side gather, for loaders that keep every frame in page-locked memory: frame i is
 * copied to points_dev + 4 * offset_i by its own cudaMemcpyAsync on `stream` (offsets = running sum of n_points_host),
 * the n_frames + 1 offsets are written to the page-locked frame_offsets_host and copied to frame_offsets_dev last.
 * Only enqueues; the frame arrays and frame_offsets_host must stay untouched until the stream has passed the copies. */
int vn_upload_frames(const float* const* frames_host, const int64_t* n_points_host, int32_t n_frames,
                     float* points_dev, int64_t capacity_points, int32_t* frame_offsets_host,
                     int32_t* frame_offsets_dev, vn_stream_t stream);

/* Number of kernel launches the library has enqueued since load (for bench.py's gpu_launches). */
int64_t vn_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* VOXELNET_B200_H */
