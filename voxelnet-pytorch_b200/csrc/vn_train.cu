// Training-mode BatchNorm support for the hot path: batch statistics and their application.
//
// The reference trains with model.train(True) (voxelnet/train.py:148), so
//   * VFELayer's BatchNorm1d (voxelnet/model.py:72,76) normalises with the mean / biased variance of
//     relu(Linear(x)) over ALL sum(K)*T rows of the batch, the zero-point pad rows included, and
//   * ConvMD's BatchNorm3d (model.py:153,161-162) with the statistics of the raw convolution output over
//     B*D*H*W positions.
// Statistics are accumulated here (fp32 partials per thread, fp64 across threads); the fold into a
// per-channel (scale, shift) and the running-statistics update are a few 16/64-element host-side
// tensor operations (voxelnet-pytorch_b200/model.py).  The convolution accumulates its own statistics in
// its epilogue (vn_conv3d_raw_stats_fwd); k_bn_relu_apply then turns the stored raw output into the next
// layer's input.
#include <cuda_bf16.h>

#include "vn_svfe_core.cuh"

namespace vn {
namespace {

constexpr int kMaxT = 64;
constexpr int kWarps = 8;

// ---- y = relu(raw * scale[c] + shift[c]) over an NDHWC bf16 tensor with 64 channels ---------------------
// out_mode 0: bf16 [P,64] (may alias raw).  One thread = 8 channels of one position (16 bytes).
__global__ void __launch_bounds__(256)
    k_bn_relu_apply_bf16(const uint4* __restrict__ raw, int64_t n_vec, const float* __restrict__ scale,
                         const float* __restrict__ shift, uint4* __restrict__ out) {
  __shared__ float s_scale[64], s_shift[64];
  if (threadIdx.x < 64) {
    s_scale[threadIdx.x] = scale[threadIdx.x];
    s_shift[threadIdx.x] = shift[threadIdx.x];
  }
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
    const uint4 v = raw[i];
    const int c0 = (int)(i & 7) * 8;
    const uint32_t in[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&in[e]);
      const float2 f = __bfloat1622float2(h);
      const float a = fmaxf(fmaf(f.x, s_scale[c0 + 2 * e], s_shift[c0 + 2 * e]), 0.0f);
      const float b = fmaxf(fmaf(f.y, s_scale[c0 + 2 * e + 1], s_shift[c0 + 2 * e + 1]), 0.0f);
      const __nv_bfloat162 r = __floats2bfloat162_rn(a, b);
      o[e] = *reinterpret_cast<const uint32_t*>(&r);
    }
    out[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// out_mode 1: f32 [B, 64*D, H, W] with channel = c*D + d (the reshape of model.py:262).
// One block iteration = kFinW consecutive w positions of one (b, d, h) row: bf16 [kFinW][64] in as 16-byte vectors
// (8 channels of one position per thread), transposed through shared memory, 64 rows of kFinW floats out as float4.
constexpr int kFinW = 64;  // w positions per block iteration: 256-byte output rows
__global__ void __launch_bounds__(256)
    k_bn_relu_apply_final(const __nv_bfloat16* __restrict__ raw, int B, int D, int H, int W,
                          const float* __restrict__ scale, const float* __restrict__ shift, float* __restrict__ out) {
  __shared__ float tile[64][kFinW + 1];
  const int p0 = threadIdx.x >> 3, c0 = (threadIdx.x & 7) * 8;
  float s[8], t[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s[j] = scale[c0 + j];
    t[j] = shift[c0 + j];
  }
  const int wblocks = (W + kFinW - 1) / kFinW;
  const int64_t n_blocks = (int64_t)B * D * H * wblocks;
  for (int64_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
    const int wb = (int)(blk % wblocks);
    int64_t r = blk / wblocks;
    const int h = (int)(r % H); r /= H;
    const int d = (int)(r % D);
    const int b = (int)(r / D);
    const int w0 = wb * kFinW;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kFinW / 32; ++i) {
      const int p = p0 + 32 * i;
      if (w0 + p < W) {
        const uint4 rv = *reinterpret_cast<const uint4*>(raw + ((((int64_t)b * D + d) * H + h) * W + w0 + p) * 64 + c0);
        const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[e]));
          tile[c0 + 2 * e][p] = fmaxf(fmaf(x.x, s[2 * e], t[2 * e]), 0.0f);
          tile[c0 + 2 * e + 1][p] = fmaxf(fmaf(x.y, s[2 * e + 1], t[2 * e + 1]), 0.0f);
        }
      }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kFinW / 16; ++i) {
      const int e = threadIdx.x + 256 * i, c = e / (kFinW / 4), q = (e % (kFinW / 4)) * 4;
      if (w0 + q < W)
        *reinterpret_cast<float4*>(out + (((int64_t)b * 64 * D + (int64_t)c * D + d) * H + h) * W + w0 + q) =
            make_float4(tile[c][q], tile[c][q + 1], tile[c][q + 2], tile[c][q + 3]);
    }
  }
}

// ---- batch statistics of the two VFE layers ----------------------------------------------------------------
// One warp per voxel on the materialised [K,T,7] feature buffer.  Rows at the tail of a voxel that are
// bit-identical to its last row (the pad rows) are evaluated once and counted with their multiplicity.
// STAGE 1: sums[0..15] += h1, sums[16..31] += h1^2 with h1 = relu(W1 x + b1)            (16 units)
// STAGE 2: VFE-1 is applied with its folded batch-statistics BN (s1, t1); the masked concat feeds VFE-2;
//          sums[0..63] += h2, sums[64..127] += h2^2 with h2 = relu(W2 in2 + b2)          (64 units)
template <int STAGE>
__global__ void __launch_bounds__(kWarps * 32, 3)
    k_svfe_stats(const float* __restrict__ feature, int64_t K, int T, const int32_t* __restrict__ n_voxels_dev,
                 const float* __restrict__ w1, const float* __restrict__ b1, const float* __restrict__ s1,
                 const float* __restrict__ t1, const float* __restrict__ w2, const float* __restrict__ b2,
                 double* __restrict__ sums) {
  extern __shared__ __align__(16) float dyn_smem[];
  __shared__ Vfe1Params p1;
  __shared__ Vfe2Shared sh;
  __shared__ double blk[128];
  const int in_ld = (T * 7 + 3) & ~3;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 16 * 7; i += blockDim.x) p1.w[i] = w1[i];
  if (threadIdx.x < 16) {
    p1.b[threadIdx.x] = b1[threadIdx.x];
    p1.s[threadIdx.x] = STAGE == 2 ? s1[threadIdx.x] : 1.0f;
    p1.t[threadIdx.x] = STAGE == 2 ? t1[threadIdx.x] : 0.0f;
  }
  if (threadIdx.x < 128) blk[threadIdx.x] = 0.0;
  Vfe2Lane q;
  if (STAGE == 2) {
    const float one = 1.0f, zero = 0.0f;
    (void)one; (void)zero;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      q.wa[k] = w2[lane * 32 + k];
      q.wb[k] = w2[(lane + 32) * 32 + k];
    }
    q.ba = b2[lane];
    q.bb = b2[lane + 32];
    for (int i = threadIdx.x; i < 16 * 64; i += blockDim.x) sh.whi[i >> 6][i & 63] = w2[(i & 63) * 32 + 16 + (i >> 6)];
  }
  __syncthreads();

  int64_t total = K;
  if (n_voxels_dev != nullptr) {
    const int64_t live = *n_voxels_dev;
    if (live < total) total = live;
  }
  float* x1 = dyn_smem + (size_t)warp * (T + 1) * kXS;
  float* in = dyn_smem + (size_t)kWarps * (T + 1) * kXS + (size_t)warp * in_ld;
  float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};  // STAGE 1: {sum, sumsq} of unit lane & 15; STAGE 2: a / b units
  const int64_t nwarps = (int64_t)gridDim.x * kWarps;
  for (int64_t v = (int64_t)blockIdx.x * kWarps + warp; v < total; v += nwarps) {
    const float* src = feature + v * (int64_t)(T * 7);
    for (int e = lane; e < T * 7; e += 32) in[e] = src[e];
    __syncwarp();
    int nu = 0;  // 1 + index of the last row that differs from row T-1
    for (int t0 = 0; t0 < T; t0 += 32) {
      const int t = t0 + lane;
      bool differs = false;
      if (t < T) {
#pragma unroll
        for (int j = 0; j < 7; ++j)
          differs |= __float_as_uint(in[t * 7 + j]) != __float_as_uint(in[(T - 1) * 7 + j]);
      }
      const uint32_t bm = __ballot_sync(0xffffffffu, differs);
      if (bm) nu = t0 + 32 - __clz(bm);
    }
    const int R = nu < T ? nu + 1 : T;      // evaluated rows: 0..nu-1, then the tail row (source T-1)
    const float tail_mult = (float)(T - nu);  // multiplicity of evaluated row nu (when nu < T)
    // VFE-1 pre-BN (STAGE 1) or post-BN (STAGE 2) values for every evaluated row
    for (int e = lane; e < R * 16; e += 32) {
      const int r = e >> 4, u = e & 15;
      const float* f = in + (r < nu ? r : T - 1) * 7;
      float a = p1.b[u];
#pragma unroll
      for (int j = 0; j < 7; ++j) a = fmaf(p1.w[u * 7 + j], f[j], a);
      const float h = fmaxf(a, 0.0f);
      if (STAGE == 1) {
        const float m = (r == nu) ? tail_mult : 1.0f;
        acc[0] = fmaf(m, h, acc[0]);
        acc[1] = fmaf(m * h, h, acc[1]);
      } else {
        x1[r * kXS + u] = fmaf(h, p1.s[u], p1.t[u]);
      }
    }
    if (STAGE == 2) {
      uint64_t on_bits = 0;
      for (int r0 = 0; r0 < R; r0 += 32) {
        const int r = r0 + lane;
        bool on = false;
        if (r < R) {
          const float* f = in + (r < nu ? r : T - 1) * 7;
          float mx = f[0];
#pragma unroll
          for (int j = 1; j < 7; ++j) mx = fmaxf(mx, f[j]);
          on = mx != 0.0f;
        }
        on_bits |= (uint64_t)__ballot_sync(0xffffffffu, on) << r0;
      }
      __syncwarp();
      if (lane < 16) {
        float a = x1[lane];
        for (int r = 1; r < R; ++r) a = fmaxf(a, x1[r * kXS + lane]);
        x1[R * kXS + lane] = a;
      }
      __syncwarp();
      // per-voxel half of VFE-2: bias + W2[:,16:32] agg1
      float ca = q.ba, cb = q.bb;
      const float* agg = x1 + R * kXS;
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        ca = fmaf(sh.whi[k][lane], agg[k], ca);
        cb = fmaf(sh.whi[k][lane + 32], agg[k], cb);
      }
      const float off_a = fmaxf(q.ba, 0.0f), off_b = fmaxf(q.bb, 0.0f);  // rows with mask 0: all-zero input
      for (int r = 0; r < R; ++r) {
        const float m = (r == nu) ? tail_mult : 1.0f;
        float ha = off_a, hb = off_b;
        if ((on_bits >> r) & 1) {
          float a = ca, b = cb;
          const float* xr = x1 + r * kXS;
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            a = fmaf(q.wa[k], xr[k], a);
            b = fmaf(q.wb[k], xr[k], b);
          }
          ha = fmaxf(a, 0.0f);
          hb = fmaxf(b, 0.0f);
        }
        acc[0] = fmaf(m, ha, acc[0]);
        acc[1] = fmaf(m * ha, ha, acc[1]);
        acc[2] = fmaf(m, hb, acc[2]);
        acc[3] = fmaf(m * hb, hb, acc[3]);
      }
    }
    __syncwarp();
  }
  if (STAGE == 1) {
    // lanes l and l + 16 hold partials of unit l & 15
    acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], 16);
    acc[1] += __shfl_xor_sync(0xffffffffu, acc[1], 16);
    if (lane < 16) {
      atomicAdd(&blk[lane], (double)acc[0]);
      atomicAdd(&blk[16 + lane], (double)acc[1]);
    }
  } else {
    atomicAdd(&blk[lane], (double)acc[0]);
    atomicAdd(&blk[64 + lane], (double)acc[1]);
    atomicAdd(&blk[32 + lane], (double)acc[2]);
    atomicAdd(&blk[96 + lane], (double)acc[3]);
  }
  __syncthreads();
  const int n_out = STAGE == 1 ? 32 : 128;
  if ((int)threadIdx.x < n_out) atomicAdd(&sums[threadIdx.x], blk[threadIdx.x]);
}

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

int vn_bn_relu_apply(const void* raw_bf16, int32_t B, int32_t D, int32_t H, int32_t W, const float* scale,
                     const float* shift, int32_t out_mode, void* out, vn_stream_t stream) {
  VN_REQUIRE(raw_bf16 && scale && shift && out, "bn_relu_apply: null pointer");
  VN_REQUIRE(B >= 1 && D >= 1 && H >= 1 && W >= 1, "bn_relu_apply: empty tensor");
  VN_REQUIRE(out_mode == 0 || out_mode == 1, "bn_relu_apply: out_mode=%d", out_mode);
  VN_REQUIRE(out_mode == 0 || W % 4 == 0, "bn_relu_apply: W=%d must be a multiple of 4 for the (B, c*D+d, H, W) layout", W);
  VN_REQUIRE((reinterpret_cast<uintptr_t>(raw_bf16) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
             "bn_relu_apply: buffers must be 16-byte aligned");
  cudaStream_t st = as_stream(stream);
  const int64_t P = (int64_t)B * D * H * W;
  if (out_mode == 0) {
    const int64_t n_vec = P * 8;
    const int64_t want = (n_vec + 255) / 256;
    const int64_t cap = (int64_t)sm_count() * 16;
    k_bn_relu_apply_bf16<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(
        reinterpret_cast<const uint4*>(raw_bf16), n_vec, scale, shift, reinterpret_cast<uint4*>(out));
    VN_LAUNCH_CHECK("k_bn_relu_apply_bf16");
  } else {
    const int64_t want = (int64_t)B * D * H * ((W + kFinW - 1) / kFinW);
    const int64_t cap = (int64_t)sm_count() * 16;
    k_bn_relu_apply_final<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(raw_bf16), B, D, H, W, scale, shift, reinterpret_cast<float*>(out));
    VN_LAUNCH_CHECK("k_bn_relu_apply_final");
  }
  return VN_OK;
}

int vn_svfe_batch_stats(const float* feature, int64_t K, int32_t T, const int32_t* n_voxels_dev, int32_t stage,
                        const float* w1, const float* b1, const float* s1, const float* t1, const float* w2,
                        const float* b2, double* sums, vn_stream_t stream) {
  VN_REQUIRE(K >= 0, "svfe_batch_stats: K < 0");
  VN_REQUIRE(T >= 1 && T <= kMaxT, "svfe_batch_stats: T=%d outside [1,%d]", T, kMaxT);
  VN_REQUIRE(stage == 1 || stage == 2, "svfe_batch_stats: stage=%d", stage);
  if (K == 0) return VN_OK;
  VN_REQUIRE(feature && w1 && b1 && sums, "svfe_batch_stats: null pointer");
  VN_REQUIRE(stage == 1 || (s1 && t1 && w2 && b2), "svfe_batch_stats: stage 2 needs the folded VFE-1 BN and the VFE-2 weights");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(sums) & 7) == 0, "svfe_batch_stats: sums must be 8-byte aligned");
  cudaStream_t st = as_stream(stream);
  const size_t smem = sizeof(float) * ((size_t)kWarps * (T + 1) * kXS + (size_t)kWarps * ((T * 7 + 3) & ~3));
  const int64_t want = (K + kWarps - 1) / kWarps;
  const int64_t cap = (int64_t)sm_count() * 3;
  const unsigned grid = (unsigned)(want < cap ? want : cap);
  if (stage == 1) {
    VN_CUDA(cudaFuncSetAttribute(k_svfe_stats<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_svfe_stats<1><<<grid, kWarps * 32, smem, st>>>(feature, K, T, n_voxels_dev, w1, b1, s1, t1, w2, b2, sums);
  } else {
    VN_CUDA(cudaFuncSetAttribute(k_svfe_stats<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_svfe_stats<2><<<grid, kWarps * 32, smem, st>>>(feature, K, T, n_voxels_dev, w1, b1, s1, t1, w2, b2, sums);
  }
  VN_LAUNCH_CHECK("k_svfe_stats");
  return VN_OK;
}

}  // extern "C"
