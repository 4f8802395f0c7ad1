// Stacked voxel feature encoder kernels: the replacement for VFELayer.forward (reference
// voxelnet/model.py:74-82) and for the VFE part of FeatureLearningNet.forward (model.py:95-100).
//
// k_svfe_fused keeps one voxel per warp entirely on chip: the [T,7] block is staged in shared
// memory, rows at the tail of a voxel that are bit-identical to its last row (the reference's pad
// rows are all [0,0,0,0,-cx,-cy,-cz]) are evaluated once (exact: max() ignores multiplicity), and
// the per-voxel arithmetic is the shared svfe_eval_voxel() of vn_svfe_core.cuh.  Only the 128
// voxel-wise maxima are written.  k_vfe_layer is the drop-in per-layer kernel.
#include "vn_svfe_core.cuh"

namespace vn {
namespace {

constexpr int kMaxT = 64;
constexpr int kWarps = 8;

// 3 blocks of 8 warps per SM: the per-row half of the VFE-2 weights lives in registers (32 per lane).
__global__ void __launch_bounds__(kWarps * 32, 3)
    k_svfe_fused(const float* __restrict__ feature, int64_t K, int T, const int32_t* __restrict__ n_voxels_dev,
                 const float* __restrict__ w1, const float* __restrict__ b1, const float* __restrict__ s1,
                 const float* __restrict__ t1, const float* __restrict__ w2, const float* __restrict__ b2,
                 const float* __restrict__ s2, const float* __restrict__ t2, float* __restrict__ voxelwise) {
  // dynamic shared memory: [kWarps][T+1][kXS] x1 rows (+ agg1) | [kWarps][in_ld] staged inputs
  extern __shared__ __align__(16) float dyn_smem[];
  __shared__ Vfe1Params p1;
  __shared__ Vfe2Shared sh;
  const int in_ld = (T * 7 + 3) & ~3;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  load_vfe1_params(p1, w1, b1, s1, t1);
  Vfe2Lane q;
  load_vfe2(q, sh, w2, b2, s2, t2, lane);
  __syncthreads();

  int64_t total = K;
  if (n_voxels_dev != nullptr) {
    const int64_t live = *n_voxels_dev;
    if (live < total) total = live;
  }
  float* x1 = dyn_smem + (size_t)warp * (T + 1) * kXS;
  float* in = dyn_smem + (size_t)kWarps * (T + 1) * kXS + (size_t)warp * in_ld;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t v = (int64_t)blockIdx.x * kWarps + warp; v < total; v += nwarps) {
    const float* src = feature + v * (int64_t)(T * 7);
    for (int e = lane; e < T * 7; e += 32) in[e] = src[e];
    __syncwarp();
    // Unique-row analysis: rows [nu, T) are bit-identical to row T-1.
    int nu = 0;  // 1 + index of the last row that differs from row T-1
    for (int t0 = 0; t0 < T; t0 += 32) {
      const int t = t0 + lane;
      bool differs = false;
      if (t < T) {
#pragma unroll
        for (int j = 0; j < 7; ++j)
          differs |= __float_as_uint(in[t * 7 + j]) != __float_as_uint(in[(T - 1) * 7 + j]);
      }
      const uint32_t b = __ballot_sync(0xffffffffu, differs);
      if (b) nu = t0 + 32 - __clz(b);
    }
    const int R = nu < T ? nu + 1 : T;  // rows to evaluate: 0..nu-1 and (if any) the tail row T-1
    float o[4];
    svfe_eval_voxel(in, nu, R, T - 1, x1, p1, q, sh, o, lane);
    store_voxel_out(o, lane, voxelwise + v * 128, nullptr);
    __syncwarp();
  }
}

// Drop-in single layer: inputs [K,T,CIN], mask [K,T] -> out [K,T,2*UNITS].  One warp per voxel;
// the voxel's rows and the layer's weights sit in shared memory (row stride CIN+1 to spread banks).
template <int CIN, int UNITS>
__global__ void __launch_bounds__(kWarps * 32)
    k_vfe_layer(const float* __restrict__ inputs, const uint8_t* __restrict__ mask, int64_t K, int T,
                const float* __restrict__ weight, const float* __restrict__ bias,
                const float* __restrict__ bn_scale, const float* __restrict__ bn_shift, float* __restrict__ out) {
  constexpr int LD = CIN + 1;
  extern __shared__ float smem[];
  float* s_w = smem;                        // [UNITS][LD]
  float* s_b = s_w + UNITS * LD;            // [UNITS]
  float* s_s = s_b + UNITS;                 // [UNITS]
  float* s_t = s_s + UNITS;                 // [UNITS]
  float* s_rows = s_t + UNITS;              // [kWarps][kMaxT][LD]
  float* s_agg = s_rows + kWarps * kMaxT * LD;  // [kWarps][UNITS]
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < UNITS * CIN; i += blockDim.x) s_w[(i / CIN) * LD + i % CIN] = weight[i];
  for (int i = threadIdx.x; i < UNITS; i += blockDim.x) {
    s_b[i] = bias[i];
    s_s[i] = bn_scale[i];
    s_t[i] = bn_shift[i];
  }
  __syncthreads();
  float* rows = s_rows + warp * kMaxT * LD;
  float* agg = s_agg + warp * UNITS;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t v = (int64_t)blockIdx.x * kWarps + warp; v < K; v += nwarps) {
    const float* src = inputs + v * (int64_t)T * CIN;
    for (int e = lane; e < T * CIN; e += 32) rows[(e / CIN) * LD + e % CIN] = src[e];
    __syncwarp();
    float* dst = out + v * (int64_t)T * (2 * UNITS);
    const uint8_t* mk = mask + v * (int64_t)T;
    // pointwise part, kept unmasked in registers' place: written masked, max taken unmasked
    for (int u0 = 0; u0 < UNITS; u0 += 32) {
      const int u = u0 + (lane % (UNITS < 32 ? UNITS : 32));
      // with UNITS == 16 the two half-warps split the rows
      const int tstep = UNITS < 32 ? 32 / UNITS : 1;
      const int tfirst = UNITS < 32 ? lane / UNITS : 0;
      float mx = -INFINITY;
      for (int t = tfirst; t < T; t += tstep) {
        float a = s_b[u];
#pragma unroll
        for (int k = 0; k < CIN; ++k) a = fmaf(s_w[u * LD + k], rows[t * LD + k], a);
        a = fmaxf(a, 0.0f) * s_s[u] + s_t[u];
        mx = fmaxf(mx, a);
        dst[(int64_t)t * (2 * UNITS) + u] = mk[t] ? a : 0.0f;
      }
      if (UNITS < 32) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
      if (UNITS >= 32 || lane < UNITS) agg[u] = mx;
    }
    __syncwarp();
    for (int e = lane; e < T * UNITS; e += 32) {
      const int t = e / UNITS, u = e % UNITS;
      dst[(int64_t)t * (2 * UNITS) + UNITS + u] = mk[t] ? agg[u] : 0.0f;
    }
    __syncwarp();
  }
}

// Training-mode statistics of ONE stand-alone layer (VFELayer.forward in .train(), model.py:74-76: BatchNorm1d over
// all K*T rows of relu(Linear(x)), pad rows included): per-unit sum and sum of squares in fp64.
// One warp per voxel like k_vfe_layer; a lane owns unit u (and u+32) and adds per-voxel fp32 partials into doubles.
template <int CIN, int UNITS>
__global__ void __launch_bounds__(kWarps * 32)
    k_vfe_layer_stats(const float* __restrict__ inputs, int64_t K, int T, const float* __restrict__ weight,
                      const float* __restrict__ bias, double* __restrict__ sums /* [2][UNITS] */) {
  constexpr int LD = CIN + 1;
  constexpr int NU = UNITS < 32 ? 1 : UNITS / 32;
  extern __shared__ float smem[];
  float* s_w = smem;                // [UNITS][LD]
  float* s_b = s_w + UNITS * LD;    // [UNITS]
  float* s_rows = s_b + UNITS;      // [kWarps][kMaxT][LD]
  __shared__ double s_acc[2 * 64];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < UNITS * CIN; i += blockDim.x) s_w[(i / CIN) * LD + i % CIN] = weight[i];
  for (int i = threadIdx.x; i < UNITS; i += blockDim.x) s_b[i] = bias[i];
  for (int i = threadIdx.x; i < 2 * 64; i += blockDim.x) s_acc[i] = 0.0;
  __syncthreads();
  float* rows = s_rows + warp * kMaxT * LD;
  double acc_s[NU], acc_q[NU];
#pragma unroll
  for (int i = 0; i < NU; ++i) { acc_s[i] = 0.0; acc_q[i] = 0.0; }
  const int tstep = UNITS < 32 ? 32 / UNITS : 1;
  const int tfirst = UNITS < 32 ? lane / UNITS : 0;
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t v = (int64_t)blockIdx.x * kWarps + warp; v < K; v += nwarps) {
    const float* src = inputs + v * (int64_t)T * CIN;
    for (int e = lane; e < T * CIN; e += 32) rows[(e / CIN) * LD + e % CIN] = src[e];
    __syncwarp();
#pragma unroll
    for (int i = 0; i < NU; ++i) {
      const int u = i * 32 + (lane % (UNITS < 32 ? UNITS : 32));
      float ps = 0.0f, pq = 0.0f;
      for (int t = tfirst; t < T; t += tstep) {
        float a = s_b[u];
#pragma unroll
        for (int k = 0; k < CIN; ++k) a = fmaf(s_w[u * LD + k], rows[t * LD + k], a);
        a = fmaxf(a, 0.0f);
        ps += a;
        pq = fmaf(a, a, pq);
      }
      acc_s[i] += (double)ps;
      acc_q[i] += (double)pq;
    }
    __syncwarp();
  }
#pragma unroll
  for (int i = 0; i < NU; ++i) {
    const int u = i * 32 + (lane % (UNITS < 32 ? UNITS : 32));
    atomicAdd(&s_acc[u], acc_s[i]);
    atomicAdd(&s_acc[64 + u], acc_q[i]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < UNITS; i += blockDim.x) {
    atomicAdd(&sums[i], s_acc[i]);
    atomicAdd(&sums[UNITS + i], s_acc[64 + i]);
  }
}

template <int CIN, int UNITS>
int launch_vfe_layer_stats(const float* inputs, int64_t K, int T, const float* weight, const float* bias, double* sums,
                           cudaStream_t st) {
  constexpr int LD = CIN + 1;
  const size_t smem = sizeof(float) * (UNITS * LD + UNITS + kWarps * kMaxT * LD);
  VN_CUDA(cudaFuncSetAttribute(k_vfe_layer_stats<CIN, UNITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  VN_CUDA(cudaMemsetAsync(sums, 0, sizeof(double) * 2 * UNITS, st));
  const int64_t want = (K + kWarps - 1) / kWarps;
  const int64_t cap = (int64_t)sm_count() * 4;
  k_vfe_layer_stats<CIN, UNITS><<<(unsigned)(want < cap ? want : cap), kWarps * 32, smem, st>>>(inputs, K, T, weight, bias, sums);
  VN_LAUNCH_CHECK("k_vfe_layer_stats");
  return VN_OK;
}

__global__ void __launch_bounds__(256)
    k_feature_mask(const float* __restrict__ feature, int64_t rows, int cin, uint8_t* __restrict__ mask) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += stride) {
    const float* f = feature + r * cin;
    float mx = f[0];
    for (int j = 1; j < cin; ++j) mx = fmaxf(mx, f[j]);
    mask[r] = mx != 0.0f ? 1 : 0;
  }
}

template <int CIN, int UNITS>
int launch_vfe_layer(const float* inputs, const uint8_t* mask, int64_t K, int T, const float* weight,
                     const float* bias, const float* bn_scale, const float* bn_shift, float* out, cudaStream_t st) {
  constexpr int LD = CIN + 1;
  const size_t smem = sizeof(float) * (UNITS * LD + 3 * UNITS + kWarps * kMaxT * LD + kWarps * UNITS);
  VN_CUDA(cudaFuncSetAttribute(k_vfe_layer<CIN, UNITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t want = (K + kWarps - 1) / kWarps;
  const int64_t cap = (int64_t)sm_count() * 4;
  k_vfe_layer<CIN, UNITS><<<(unsigned)(want < cap ? want : cap), kWarps * 32, smem, st>>>(
      inputs, mask, K, T, weight, bias, bn_scale, bn_shift, out);
  VN_LAUNCH_CHECK("k_vfe_layer");
  return VN_OK;
}

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

int vn_vfe_layer_fwd(const float* inputs, const uint8_t* mask, int64_t K, int32_t T, int32_t cin, int32_t units,
                     const float* weight, const float* bias, const float* bn_scale, const float* bn_shift,
                     float* out, vn_stream_t stream) {
  VN_REQUIRE(K >= 0, "vfe_layer: K < 0");
  VN_REQUIRE(T >= 1 && T <= kMaxT, "vfe_layer: T=%d outside [1,%d]", T, kMaxT);
  if (K == 0) return VN_OK;
  VN_REQUIRE(inputs && mask && weight && bias && bn_scale && bn_shift && out, "vfe_layer: null pointer");
  cudaStream_t st = as_stream(stream);
  if (cin == 7 && units == 16) return launch_vfe_layer<7, 16>(inputs, mask, K, T, weight, bias, bn_scale, bn_shift, out, st);
  if (cin == 32 && units == 64) return launch_vfe_layer<32, 64>(inputs, mask, K, T, weight, bias, bn_scale, bn_shift, out, st);
  set_error("vfe_layer: unsupported (cin=%d, units=%d); the reference instantiates (7,16) and (32,64)", cin, units);
  return VN_ERR_UNSUPPORTED;
}

int vn_vfe_layer_batch_stats(const float* inputs, int64_t K, int32_t T, int32_t cin, int32_t units, const float* weight,
                             const float* bias, double* sums, vn_stream_t stream) {
  VN_REQUIRE(K >= 1, "vfe_layer_batch_stats: K < 1 (BatchNorm needs at least one row)");
  VN_REQUIRE(T >= 1 && T <= kMaxT, "vfe_layer_batch_stats: T=%d outside [1,%d]", T, kMaxT);
  VN_REQUIRE(inputs && weight && bias && sums, "vfe_layer_batch_stats: null pointer");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(sums) & 7) == 0, "vfe_layer_batch_stats: sums must be 8-byte aligned");
  cudaStream_t st = as_stream(stream);
  if (cin == 7 && units == 16) return launch_vfe_layer_stats<7, 16>(inputs, K, T, weight, bias, sums, st);
  if (cin == 32 && units == 64) return launch_vfe_layer_stats<32, 64>(inputs, K, T, weight, bias, sums, st);
  set_error("vfe_layer_batch_stats: unsupported (cin=%d, units=%d); the reference instantiates (7,16) and (32,64)", cin, units);
  return VN_ERR_UNSUPPORTED;
}

int vn_feature_mask(const float* feature, int64_t K, int32_t T, int32_t cin, uint8_t* mask, vn_stream_t stream) {
  VN_REQUIRE(K >= 0 && T >= 1 && cin >= 1, "feature_mask: bad shape");
  if (K == 0) return VN_OK;
  VN_REQUIRE(feature && mask, "feature_mask: null pointer");
  const int64_t rows = K * T;
  const int64_t want = (rows + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  k_feature_mask<<<(unsigned)(want < cap ? want : cap), 256, 0, as_stream(stream)>>>(feature, rows, cin, mask);
  VN_LAUNCH_CHECK("k_feature_mask");
  return VN_OK;
}

int vn_svfe_fused_fwd(const float* feature, int64_t K, int32_t T, const int32_t* n_voxels_dev, const float* w1,
                      const float* b1, const float* s1, const float* t1, const float* w2, const float* b2,
                      const float* s2, const float* t2, float* voxelwise, vn_stream_t stream) {
  VN_REQUIRE(K >= 0, "svfe_fused: K < 0");
  VN_REQUIRE(T >= 1 && T <= kMaxT, "svfe_fused: T=%d outside [1,%d]", T, kMaxT);
  if (K == 0) return VN_OK;
  VN_REQUIRE(feature && w1 && b1 && s1 && t1 && w2 && b2 && s2 && t2 && voxelwise, "svfe_fused: null pointer");
  cudaStream_t st = as_stream(stream);
  const size_t smem = sizeof(float) * ((size_t)kWarps * (T + 1) * kXS + (size_t)kWarps * ((T * 7 + 3) & ~3));
  VN_CUDA(cudaFuncSetAttribute(k_svfe_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t want = (K + kWarps - 1) / kWarps;
  const int64_t cap = (int64_t)sm_count() * 3;
  k_svfe_fused<<<(unsigned)(want < cap ? want : cap), kWarps * 32, smem, st>>>(feature, K, T, n_voxels_dev, w1, b1, s1,
                                                                              t1, w2, b2, s2, t2, voxelwise);
  VN_LAUNCH_CHECK("k_svfe_fused");
  return VN_OK;
}

}  // extern "C"
