// Sparse -> dense scatter: the replacement for
//   torch.sparse.FloatTensor(coordinate.t(), voxelwise, [B, D, H, W, 128]).to_dense()
// (reference voxelnet/model.py:102-106).  `dense` is [B, D, H, W, C] contiguous (NDHWC), which is
// also the physical layout the Conv3d middle layers consume after the reference's permute
// (model.py:259 produces a channels-last view).
//
// One warp moves one voxel row (C = 128 floats = 512 B in, 512 B fp32 or 256 B bf16 out), fully
// coalesced on both sides; the only irregular access is the row's base address.
#include <cuda_bf16.h>

#include "vn_common.cuh"

namespace vn {
namespace {

__device__ __forceinline__ bool cell_offset(const long long* c, int B, int D, int H, int W, int64_t& cell) {
  const long long b = c[0], z = c[1], y = c[2], x = c[3];
  if (b < 0 || b >= B || z < 0 || z >= D || y < 0 || y >= H || x < 0 || x >= W) return false;
  cell = ((b * D + z) * H + y) * (int64_t)W + x;
  return true;
}

__device__ __forceinline__ int64_t live_rows(int64_t K, const int32_t* n_voxels_dev) {
  if (n_voxels_dev != nullptr) {
    const int64_t live = *n_voxels_dev;
    if (live < K) K = live;
  }
  return K;
}

// fp32 drop-in.  COO to_dense() sums duplicate coordinates, so rows are accumulated with
// red.global.add.f32 (a plain store when coordinates are unique, as the voxelizer guarantees).
__global__ void __launch_bounds__(256)
    k_scatter_f32(const float* __restrict__ voxelwise, const long long* __restrict__ coords, int64_t K,
                  const int32_t* __restrict__ n_voxels_dev, int C, int B, int D, int H, int W,
                  float* __restrict__ dense) {
  const int64_t total = live_rows(K, n_voxels_dev);
  const uint32_t lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < total; r += nwarps) {
    int64_t cell;
    if (!cell_offset(coords + r * 4, B, D, H, W, cell)) continue;
    const float* src = voxelwise + r * C;
    float* dst = dense + cell * C;
    for (int c = lane; c < C; c += 32) atomicAdd(dst + c, src[c]);
  }
}

// bf16 fast path: convert on the fly (round-to-nearest-even) and store 8 B per lane per step, or
// write zeros (clear != 0) to return the resident grid to all-zero after the convolution read it.
__global__ void __launch_bounds__(256)
    k_scatter_bf16(const float* __restrict__ voxelwise, const long long* __restrict__ coords, int64_t K,
                   const int32_t* __restrict__ n_voxels_dev, int C, int B, int D, int H, int W,
                   __nv_bfloat16* __restrict__ dense, int clear) {
  const int64_t total = live_rows(K, n_voxels_dev);
  const uint32_t lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp; r < total; r += nwarps) {
    int64_t cell;
    if (!cell_offset(coords + r * 4, B, D, H, W, cell)) continue;
    uint2* dst = reinterpret_cast<uint2*>(dense + cell * C);
    if (clear) {
      for (int c4 = lane; c4 < C / 4; c4 += 32) dst[c4] = make_uint2(0u, 0u);
    } else {
      const float4* src = reinterpret_cast<const float4*>(voxelwise + r * C);
      for (int c4 = lane; c4 < C / 4; c4 += 32) {
        const float4 v = src[c4];
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
        const __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
        dst[c4] = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
      }
    }
  }
}

int check(const void* voxelwise, const int64_t* coords, int64_t K, int C, int B, int D, int H, int W,
          const void* dense, bool need_src) {
  VN_REQUIRE(K >= 0, "scatter: K < 0");
  VN_REQUIRE(C > 0 && C % 4 == 0, "scatter: C=%d must be a positive multiple of 4", C);
  VN_REQUIRE(B > 0 && D > 0 && H > 0 && W > 0, "scatter: empty grid");
  if (K == 0) return VN_OK;
  VN_REQUIRE(coords != nullptr && dense != nullptr, "scatter: null pointer");
  VN_REQUIRE(!need_src || voxelwise != nullptr, "scatter: voxelwise is null");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(voxelwise) & 15) == 0 && (reinterpret_cast<uintptr_t>(dense) & 15) == 0,
             "scatter: buffers must be 16-byte aligned");
  return VN_OK;
}

unsigned grid_for_rows(int64_t K) {
  const int64_t want = (K + 7) / 8;  // 8 warps per block
  const int64_t cap = (int64_t)sm_count() * 8;
  return (unsigned)(want < cap ? (want > 0 ? want : 1) : cap);
}

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

int vn_scatter_dense_f32(const float* voxelwise, const int64_t* coords, int64_t K, const int32_t* n_voxels_dev,
                         int32_t C, int32_t B, int32_t D, int32_t H, int32_t W, float* dense, vn_stream_t stream) {
  int rc = check(voxelwise, coords, K, C, B, D, H, W, dense, true);
  if (rc != VN_OK || K == 0) return rc;
  k_scatter_f32<<<grid_for_rows(K), 256, 0, as_stream(stream)>>>(
      voxelwise, reinterpret_cast<const long long*>(coords), K, n_voxels_dev, C, B, D, H, W, dense);
  VN_LAUNCH_CHECK("k_scatter_f32");
  return VN_OK;
}

int vn_scatter_dense_bf16(const float* voxelwise, const int64_t* coords, int64_t K, const int32_t* n_voxels_dev,
                          int32_t C, int32_t B, int32_t D, int32_t H, int32_t W, void* dense_bf16, int32_t clear,
                          vn_stream_t stream) {
  int rc = check(voxelwise, coords, K, C, B, D, H, W, dense_bf16, clear == 0);
  if (rc != VN_OK || K == 0) return rc;
  k_scatter_bf16<<<grid_for_rows(K), 256, 0, as_stream(stream)>>>(
      voxelwise, reinterpret_cast<const long long*>(coords), K, n_voxels_dev, C, B, D, H, W,
      reinterpret_cast<__nv_bfloat16*>(dense_bf16), clear);
  VN_LAUNCH_CHECK("k_scatter_bf16");
  return VN_OK;
}

}  // extern "C"
