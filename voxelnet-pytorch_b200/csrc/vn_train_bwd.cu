// Backward of the training-mode hot path, everything except the two dense conv GEMMs (vn_conv3d_bwd.cu):
//
//   * BatchNorm3d(batch statistics) + ReLU backward of a ConvMD layer (voxelnet/model.py:161-165; the reference
//     gets it from autograd, loss.backward() at train.py:149):
//         dz = dy * (relu input > 0),  dgamma = sum dz*xhat,  dbeta = sum dz,
//         draw = gamma*invstd * (dz - dbeta/N - xhat*dgamma/N)
//     as a reduction pass and an apply pass over the stored raw convolution output.  The last layer reads dy
//     in the (B, c*D+d, H, W) fp32 layout of the reshape at model.py:262.
//   * data gradient of the FIRST conv layer evaluated only at the occupied voxel cells and written as the
//     gradient of the (K,128) voxel-wise features: backward of to_dense (model.py:102-106) fused with the
//     transposed convolution - the 0.7 GB/frame dense gradient grid never exists.
//   * backward of the stacked VFE (model.py:74-82, 95-100) with batch-statistics BatchNorm1d over all K*T rows
//     (pad rows included): three passes over the voxels, each recomputing the forward on chip.
#include <cuda_bf16.h>

#include "vn_common.cuh"

namespace vn {
namespace {

// ---------------------------------------------------------------------------------------------------------
// BatchNorm3d + ReLU backward.  prm: [6][64] = scale (gamma*invstd), shift, mean, invstd, c1 = dbeta/N,
// c2 = dgamma/N (rows 4, 5 only read by the apply pass).
// ---------------------------------------------------------------------------------------------------------
template <bool APPLY>
__global__ void __launch_bounds__(256)
    k_bn_bwd_ndhwc(const uint4* __restrict__ raw, const uint4* __restrict__ dy, int64_t n_vec,
                   const float* __restrict__ prm, double* __restrict__ sums, uint4* __restrict__ draw) {
  __shared__ float blk[128];
  if (!APPLY) {
    if (threadIdx.x < 128) blk[threadIdx.x] = 0.0f;
    __syncthreads();
  }
  const int c0 = (int)(threadIdx.x & 7) * 8;  // blockDim and the grid stride are multiples of 8
  float s[8], t[8], mu[8], is[8], k1[8], k2[8], a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s[j] = prm[c0 + j];
    t[j] = prm[64 + c0 + j];
    mu[j] = prm[128 + c0 + j];
    is[j] = prm[192 + c0 + j];
    k1[j] = APPLY ? prm[256 + c0 + j] : 0.0f;
    k2[j] = APPLY ? prm[320 + c0 + j] : 0.0f;
    a1[j] = 0.0f;
    a2[j] = 0.0f;
  }
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
    const uint4 rv = raw[i], gv = dy[i];
    const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w}, gw[4] = {gv.x, gv.y, gv.z, gv.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[e]));
      const float2 g = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&gw[e]));
      const float xs[2] = {x.x, x.y}, gs[2] = {g.x, g.y};
      float d[2];
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int j = 2 * e + h;
        const float dz = fmaf(xs[h], s[j], t[j]) > 0.0f ? gs[h] : 0.0f;
        const float xh = (xs[h] - mu[j]) * is[j];
        if (APPLY) {
          d[h] = s[j] * (dz - k1[j] - xh * k2[j]);
        } else {
          a1[j] += dz;
          a2[j] = fmaf(dz, xh, a2[j]);
        }
      }
      if (APPLY) {
        const __nv_bfloat162 r = __floats2bfloat162_rn(d[0], d[1]);
        o[e] = *reinterpret_cast<const uint32_t*>(&r);
      }
    }
    if (APPLY) draw[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
  if (!APPLY) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&blk[c0 + j], a1[j]);
      atomicAdd(&blk[64 + c0 + j], a2[j]);
    }
    __syncthreads();
    if (threadIdx.x < 128) atomicAdd(&sums[threadIdx.x], (double)blk[threadIdx.x]);
  }
}

// dy in the final layout: f32 [B, 64*D, H, W], channel = c*D + d.  One block iteration = kFinW consecutive w of one
// (b,d,h): the 64 x kFinW fp32 dy tile is read as 256-byte rows (float4 per thread) and transposed through shared
// memory; raw / draw move as 16-byte vectors (8 channels of one position per thread).  W % 4 == 0.
constexpr int kFinW = 64;  // w positions per block iteration: 256-byte dy rows
template <bool APPLY>
__global__ void __launch_bounds__(256)
    k_bn_bwd_final(const __nv_bfloat16* __restrict__ raw, const float* __restrict__ dy, int B, int D, int H, int W,
                   const float* __restrict__ prm, double* __restrict__ sums, __nv_bfloat16* __restrict__ draw) {
  __shared__ float tile[64][kFinW + 1];
  __shared__ float blk[128];
  if (threadIdx.x < 128) blk[threadIdx.x] = 0.0f;
  const int p0 = threadIdx.x >> 3, c0 = (threadIdx.x & 7) * 8;  // compute role: positions p0 + 32 i, channels c0 .. c0+7
  float s[8], t[8], mu[8], is[8], k1[8], k2[8], a1[8], a2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    s[j] = prm[c0 + j];
    t[j] = prm[64 + c0 + j];
    mu[j] = prm[128 + c0 + j];
    is[j] = prm[192 + c0 + j];
    k1[j] = APPLY ? prm[256 + c0 + j] : 0.0f;
    k2[j] = APPLY ? prm[320 + c0 + j] : 0.0f;
    a1[j] = 0.0f;
    a2[j] = 0.0f;
  }
  const int wblocks = (W + kFinW - 1) / kFinW;
  const int64_t n_blocks = (int64_t)B * D * H * wblocks;
  for (int64_t b_ = blockIdx.x; b_ < n_blocks; b_ += gridDim.x) {
    const int wb = (int)(b_ % wblocks);
    int64_t r = b_ / wblocks;
    const int h = (int)(r % H); r /= H;
    const int d = (int)(r % D);
    const int b = (int)(r / D);
    const int w0 = wb * kFinW;
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kFinW / 16; ++i) {  // 64 rows x (kFinW / 4) float4
      const int e = threadIdx.x + 256 * i, cc = e / (kFinW / 4), q = (e % (kFinW / 4)) * 4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (w0 + q < W)
        v = *reinterpret_cast<const float4*>(dy + (((int64_t)b * 64 * D + (int64_t)cc * D + d) * H + h) * W + w0 + q);
      tile[cc][q] = v.x; tile[cc][q + 1] = v.y; tile[cc][q + 2] = v.z; tile[cc][q + 3] = v.w;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kFinW / 32; ++i) {
      const int p = p0 + 32 * i;
      if (w0 + p < W) {
        const int64_t off = ((((int64_t)b * D + d) * H + h) * W + w0 + p) * 64 + c0;
        const uint4 rv = *reinterpret_cast<const uint4*>(raw + off);
        const uint32_t rw[4] = {rv.x, rv.y, rv.z, rv.w};
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&rw[e]));
          const float xs[2] = {x.x, x.y};
          float dd[2];
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const int j = 2 * e + hh;
            const float dz = fmaf(xs[hh], s[j], t[j]) > 0.0f ? tile[c0 + j][p] : 0.0f;
            const float xh = (xs[hh] - mu[j]) * is[j];
            if (APPLY) {
              dd[hh] = s[j] * (dz - k1[j] - xh * k2[j]);
            } else {
              a1[j] += dz;
              a2[j] = fmaf(dz, xh, a2[j]);
            }
          }
          if (APPLY) {
            const __nv_bfloat162 rr = __floats2bfloat162_rn(dd[0], dd[1]);
            o[e] = *reinterpret_cast<const uint32_t*>(&rr);
          }
        }
        if (APPLY) *reinterpret_cast<uint4*>(draw + off) = make_uint4(o[0], o[1], o[2], o[3]);
      }
    }
  }
  if (!APPLY) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      atomicAdd(&blk[c0 + j], a1[j]);
      atomicAdd(&blk[64 + c0 + j], a2[j]);
    }
    __syncthreads();
    if (threadIdx.x < 128) atomicAdd(&sums[threadIdx.x], (double)blk[threadIdx.x]);
  }
}

// ---------------------------------------------------------------------------------------------------------
// First conv layer: data gradient at the occupied cells only, as a gather-GEMM on tcgen05.
//   dvw[v][c] = sum over taps (kd,kh,kw) and n of draw[b, d', y+1-kh, x+1-kw][n] * W[n][c][kd][kh][kw],
//   d' = (z + pad - kd) / stride when divisible and inside [0, Dr).
// CTA = 128 consecutive voxels (np.unique order: they mostly share z, so the taps of the other depth parity drop
// out for the whole CTA).  Per tap: thread r gathers voxel r's 128-byte gradient row (or zeros) into row r of a
// SWIZZLE_128B K-major A tile and copies row r of the tap's [128 c][64 n] weight slice into the B tile; one
// thread then issues 4 x tcgen05.mma (M = 128 voxels, N = 128 channels, K = 16) accumulating into a 128-column
// TMEM tile.  Two stages: the gather of the next tap overlaps the MMAs of this one.
// ---------------------------------------------------------------------------------------------------------
constexpr int kGV = 128;
constexpr uint32_t kGTile = 128 * 128;  // 16 KB: 128 rows x 64 bf16

struct GatherCtl {
  uint64_t mma_done[2];
  uint32_t tmem_base;
  uint32_t pad;
};
constexpr uint32_t kGatherSmem = 4 * kGTile + 1024 + (uint32_t)sizeof(GatherCtl);

__global__ void __launch_bounds__(kGV)
    k_conv1_dgrad_gather(const __nv_bfloat16* __restrict__ draw, const __nv_bfloat16* __restrict__ wt /* [27][128][64] */,
                         const long long* __restrict__ coords, int64_t K, int Dr, int H, int W, int stride_d,
                         int pad_d, float* __restrict__ dvw) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  GatherCtl* ctl = reinterpret_cast<GatherCtl*>(smem + 4 * kGTile);
  const uint32_t r = threadIdx.x, warp = r >> 5;
  if (r == 0) {
    mbar_init(&ctl->mma_done[0], 1);
    mbar_init(&ctl->mma_done[1], 1);
    mbar_fence_init();
  }
  if (warp == 0) {
    tmem_alloc(&ctl->tmem_base, 128);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = ctl->tmem_base;

  const int64_t v = (int64_t)blockIdx.x * kGV + r;
  int vb = -1, vz = 0, vy = 0, vx = 0;
  if (v < K) {
    vb = (int)coords[v * 4 + 0];
    vz = (int)coords[v * 4 + 1];
    vy = (int)coords[v * 4 + 2];
    vx = (int)coords[v * 4 + 3];
  }
  const uint32_t row_off = r * 128, sw = r & 7;
  const uint32_t idesc = umma_idesc_bf16(128, 128);
  uint32_t it = 0;
  for (int tap = 0; tap < 27; ++tap) {
    const int kd = tap / 9, kh = (tap / 3) % 3, kw = tap % 3;
    const int zz = vz + pad_d - kd;
    const int dd = zz / stride_d;
    const int hh = vy + 1 - kh, ww = vx + 1 - kw;
    const bool valid = vb >= 0 && zz >= 0 && (zz % stride_d) == 0 && dd < Dr && hh >= 0 && hh < H && ww >= 0 && ww < W;
    if (!__syncthreads_or(valid)) continue;
    const uint32_t s = it & 1;
    if (it >= 2) mbar_wait(&ctl->mma_done[s], ((it >> 1) - 1) & 1);  // the MMAs that read this stage are done
    uint8_t* a_tile = smem + s * 2 * kGTile;
    uint8_t* b_tile = a_tile + kGTile;
    {
      uint4 q[8];
      if (valid) {
        const uint4* src = reinterpret_cast<const uint4*>(draw + ((((int64_t)vb * Dr + dd) * H + hh) * W + ww) * 64);
#pragma unroll
        for (int j = 0; j < 8; ++j) q[j] = src[j];
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) q[j] = make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(a_tile + row_off + (((uint32_t)j ^ sw) << 4)) = q[j];
      const uint4* wsrc = reinterpret_cast<const uint4*>(wt + ((size_t)tap * 128 + r) * 64);
#pragma unroll
      for (int j = 0; j < 8; ++j) q[j] = wsrc[j];
#pragma unroll
      for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(b_tile + row_off + (((uint32_t)j ^ sw) << 4)) = q[j];
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0) {
      if (elect_one()) {
        tc_fence_after_sync();
        const uint32_t a_addr = smem_u32(a_tile), b_addr = smem_u32(b_tile);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          umma_bf16_ss(tmem_base, umma_desc_sw128(a_addr + k * 32, 1024), umma_desc_sw128(b_addr + k * 32, 1024), idesc,
                       (it == 0 && k == 0) ? 0u : 1u);
        umma_commit(&ctl->mma_done[s]);
      }
      __syncwarp();
    }
    ++it;
  }
  if (it > 0) mbar_wait(&ctl->mma_done[(it - 1) & 1], ((it - 1) >> 1) & 1);
  tc_fence_after_sync();
  {  // tcgen05.ld is warp-collective: every lane loads its row, only rows of real voxels are stored
    float* dst = dvw + v * 128;
#pragma unroll
    for (int c4 = 0; c4 < 4; ++c4) {
      uint32_t o[32];
      if (it > 0) {  // block-uniform
        tmem_ld_32x32b_x32(tmem_base + ((warp * 32) << 16) + c4 * 32, o);
        tmem_ld_wait();
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = 0u;
      }
      if (v < K) {
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          *reinterpret_cast<uint4*>(dst + c4 * 32 + j) = make_uint4(o[j], o[j + 1], o[j + 2], o[j + 3]);
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, 128);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Stacked-VFE backward.  One warp per voxel on the materialised [K,T,7] feature buffer; rows at the tail of a
// voxel that are bit-identical to its last row (the pad rows) are evaluated once with their multiplicity m.
//
// Forward (recomputed):  h1 = relu(W1 f + b1), n1 = h1*s1 + t1, agg1 = max_r n1, c1 = on_r * [n1 | agg1],
//                        h2 = relu(W2 c1 + b2), n2 = h2*s2 + t2, agg2 = max_r n2 (all rows),
//                        vw[v] = max_r on_r*n2[r][v], vw[64+v] = max_r on_r*agg2[v].
// Backward: the max picks one row per (voxel, unit) - a sparse dn2 - but batch-statistics BatchNorm spreads
//           -(dbeta + xhat*dgamma)/N over EVERY row, pads included, so three passes are needed:
//   PASS 1: dbeta2, dgamma2                       (needs dvw)
//   PASS 2: dW2, db2, dbeta1, dgamma1             (needs dbeta2/N, dgamma2/N)
//   PASS 3: dW1, db1                              (needs dbeta1/N, dgamma1/N and PASS 2's row gradients)
// bn1: [6][16], bn2: [6][64] = scale, shift, mean, invstd, dbeta/N, dgamma/N.
// acc (fp64): dW2[64][32] @0, db2 @2048, dgamma2 @2112, dbeta2 @2176, dW1[16][7] @2240, db1 @2352,
//             dgamma1 @2368, dbeta1 @2384; 2400 values.
// ---------------------------------------------------------------------------------------------------------
// PASS 2 also stores each evaluated row's VFE-1 output gradient (16 floats) to `rows` [K][T][16], so that PASS 3
// only redoes the cheap VFE-1 forward.
constexpr int kBwWarps = 8;
constexpr int kAccN = 2400;
constexpr int kOffW2 = 0, kOffB2 = 2048, kOffG2 = 2112, kOffBe2 = 2176, kOffW1 = 2240, kOffB1 = 2352, kOffG1 = 2368,
              kOffBe1 = 2384;
constexpr int kW2kN = 2 * 32 * 16 + 16;  // [half][u][k] with the second half 16 floats further (bank spread)

struct SvfeBwdShared {
  float w1[16 * 7], b1[16];
  float bn1[6][16];
  float b2[64];
  float bn2[6][64];
  float whi[16][64];   // whi[k][unit] = W2[unit][16 + k]          (forward, per-voxel half)
  float w2lo[kW2kN];   // W2[unit][k],      k < 16, as [half][unit & 31][k]  (gradient of the per-row half)
  float w2hi[kW2kN];   // W2[unit][16 + k], k < 16, same layout              (gradient of agg1)
  float blk[kAccN];
};
__device__ __forceinline__ int w2k_index(int unit, int k) { return (unit >> 5) * (32 * 16 + 16) + (unit & 31) * 16 + k; }

template <int PASS>
__global__ void __launch_bounds__(kBwWarps * 32, 2)
    k_svfe_bwd(const float* __restrict__ feature, int64_t K, int T, const float* __restrict__ dvw,
               const float* __restrict__ w1, const float* __restrict__ b1, const float* __restrict__ w2,
               const float* __restrict__ b2, const float* __restrict__ bn1, const float* __restrict__ bn2,
               float* __restrict__ rows, double* __restrict__ acc_out) {
  extern __shared__ __align__(16) float dyn_smem[];
  __shared__ SvfeBwdShared S;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 16 * 7; i += blockDim.x) S.w1[i] = w1[i];
  for (int i = threadIdx.x; i < 16; i += blockDim.x) S.b1[i] = b1[i];
  for (int i = threadIdx.x; i < 6 * 16; i += blockDim.x) S.bn1[i >> 4][i & 15] = bn1[i];
  for (int i = threadIdx.x; i < kAccN; i += blockDim.x) S.blk[i] = 0.0f;
  if (PASS != 3) {
    for (int i = threadIdx.x; i < 64; i += blockDim.x) S.b2[i] = b2[i];
    for (int i = threadIdx.x; i < 6 * 64; i += blockDim.x) S.bn2[i >> 6][i & 63] = bn2[i];
    for (int i = threadIdx.x; i < 16 * 64; i += blockDim.x) S.whi[i >> 6][i & 63] = w2[(i & 63) * 32 + 16 + (i >> 6)];
    for (int i = threadIdx.x; i < 64 * 16; i += blockDim.x) {
      const int unit = i >> 4, k = i & 15;
      S.w2lo[w2k_index(unit, k)] = w2[unit * 32 + k];
      S.w2hi[w2k_index(unit, k)] = w2[unit * 32 + 16 + k];
    }
  }
  __syncthreads();

  // per-warp scratch: in[T*7 (+pad)], n1[(T+1)*16], h1[T*16], g1[T*16], dzs[72]
  const int in_ld = (T * 7 + 3) & ~3;
  const int per_warp = in_ld + (T + 1) * 16 + T * 16 + T * 16 + 72;
  float* in = dyn_smem + (size_t)warp * per_warp;
  float* n1 = in + in_ld;
  float* h1s = n1 + (T + 1) * 16;
  float* g1 = h1s + T * 16;
  float* dzs = g1 + T * 16;  // unit u at dzs[u + (u >> 5)] (the two halves in different banks)

  // lane owns VFE-2 units a = lane, b = lane + 32
  float wa[16], wb[16];
  float b2a = 0.f, b2b = 0.f, s2a = 0.f, s2b = 0.f, t2a = 0.f, t2b = 0.f, m2a = 0.f, m2b = 0.f, i2a = 0.f, i2b = 0.f;
  float A2a = 0.f, A2b = 0.f, B2a = 0.f, B2b = 0.f;
  if (PASS != 3) {
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      wa[k] = w2[lane * 32 + k];
      wb[k] = w2[(lane + 32) * 32 + k];
    }
    b2a = S.b2[lane]; b2b = S.b2[lane + 32];
    s2a = S.bn2[0][lane]; s2b = S.bn2[0][lane + 32];
    t2a = S.bn2[1][lane]; t2b = S.bn2[1][lane + 32];
    m2a = S.bn2[2][lane]; m2b = S.bn2[2][lane + 32];
    i2a = S.bn2[3][lane]; i2b = S.bn2[3][lane + 32];
    A2a = S.bn2[4][lane]; A2b = S.bn2[4][lane + 32];
    B2a = S.bn2[5][lane]; B2b = S.bn2[5][lane + 32];
  }
  const int u1 = lane & 15;  // VFE-1 unit of this lane in the (row, unit) sweeps; also the k of the dc sweeps
  const int half = lane >> 4;
  const float s1u = S.bn1[0][u1], t1u = S.bn1[1][u1], m1u = S.bn1[2][u1], i1u = S.bn1[3][u1];
  const float A1u = S.bn1[4][u1], B1u = S.bn1[5][u1];

  // accumulators
  float dbe2a = 0.f, dbe2b = 0.f, dg2a = 0.f, dg2b = 0.f;           // PASS 1
  float dWa[32], dWb[32];                                             // PASS 2: dW2[a][0..31], dW2[b][0..31]
  float db2a = 0.f, db2b = 0.f, dbe1 = 0.f, dg1 = 0.f;               // PASS 2
  float dW1[7], db1 = 0.f;                                            // PASS 3
#pragma unroll
  for (int k = 0; k < 32; ++k) { dWa[k] = 0.f; dWb[k] = 0.f; }
#pragma unroll
  for (int j = 0; j < 7; ++j) dW1[j] = 0.f;

  const int64_t nwarps = (int64_t)gridDim.x * kBwWarps;
  for (int64_t v = (int64_t)blockIdx.x * kBwWarps + warp; v < K; v += nwarps) {
    const float* src = feature + v * (int64_t)(T * 7);
    for (int e = lane; e < T * 7; e += 32) in[e] = src[e];
    __syncwarp();
    int nu = 0;  // 1 + index of the last row that differs from row T-1
    for (int t0 = 0; t0 < T; t0 += 32) {
      const int t = t0 + lane;
      bool differs = false;
      if (t < T) {
#pragma unroll
        for (int j = 0; j < 7; ++j) differs |= __float_as_uint(in[t * 7 + j]) != __float_as_uint(in[(T - 1) * 7 + j]);
      }
      const uint32_t bm = __ballot_sync(0xffffffffu, differs);
      if (bm) nu = t0 + 32 - __clz(bm);
    }
    const int R = nu < T ? nu + 1 : T;
    const float tail_mult = (float)(T - nu);
    float* grow = rows + v * (int64_t)(T * 16);
    // ---- VFE-1 forward ----
    for (int e = lane; e < R * 16; e += 32) {
      const int r = e >> 4;
      const float* f = in + (r < nu ? r : T - 1) * 7;
      float a = S.b1[u1];
#pragma unroll
      for (int j = 0; j < 7; ++j) a = fmaf(S.w1[u1 * 7 + j], f[j], a);
      const float h = fmaxf(a, 0.0f);
      h1s[e] = h;
      n1[e] = fmaf(h, s1u, t1u);
      g1[e] = PASS == 3 ? grow[e] : 0.0f;
    }
    __syncwarp();
    if (PASS == 3) {
      // ---- VFE-1 backward from the stored row gradients ----
      for (int e = lane; e < R * 16; e += 32) {
        const int r = e >> 4;
        const float g = g1[e], h = h1s[e];
        const float xh = (h - m1u) * i1u;
        const float m = (r == nu) ? tail_mult : 1.0f;
        const float dz = h > 0.0f ? s1u * (g - m * fmaf(xh, B1u, A1u)) : 0.0f;
        const float* f = in + (r < nu ? r : T - 1) * 7;
        db1 += dz;
#pragma unroll
        for (int j = 0; j < 7; ++j) dW1[j] = fmaf(dz, f[j], dW1[j]);
      }
      __syncwarp();
      continue;
    }
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
    int arg1 = 0;
    if (lane < 16) {
      float a = n1[lane];
      for (int r = 1; r < R; ++r) {
        const float x = n1[r * 16 + lane];
        if (x > a) { a = x; arg1 = r; }
      }
      n1[R * 16 + lane] = a;
    }
    __syncwarp();
    const float* agg1 = n1 + R * 16;
    const uint64_t evaluated = R >= 64 ? ~0ull : ((1ull << R) - 1ull);
    const bool any_on = on_bits != 0, any_off = on_bits != evaluated;
    // ---- VFE-2 forward, per-voxel half ----
    float ca = b2a, cb = b2b;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      ca = fmaf(S.whi[k][lane], agg1[k], ca);
      cb = fmaf(S.whi[k][lane + 32], agg1[k], cb);
    }
    // ---- sweep 1: maxima and their rows ----
    float on_a = -INFINITY, on_b = -INFINITY, all_a = -INFINITY, all_b = -INFINITY;
    float hon_a = 0.f, hon_b = 0.f, hall_a = 0.f, hall_b = 0.f;
    int ron_a = -1, ron_b = -1, rall_a = -1, rall_b = -1;
    for (int r = 0; r < R; ++r) {
      const bool on = (on_bits >> r) & 1;
      float pa = b2a, pb = b2b;
      if (on) {
        pa = ca; pb = cb;
        const float* xr = n1 + r * 16;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          pa = fmaf(wa[k], xr[k], pa);
          pb = fmaf(wb[k], xr[k], pb);
        }
      }
      const float ha = fmaxf(pa, 0.0f), hb = fmaxf(pb, 0.0f);
      const float na = fmaf(ha, s2a, t2a), nb = fmaf(hb, s2b, t2b);
      if (on && na > on_a) { on_a = na; ron_a = r; hon_a = ha; }
      if (on && nb > on_b) { on_b = nb; ron_b = r; hon_b = hb; }
      if (na > all_a) { all_a = na; rall_a = r; hall_a = ha; }
      if (nb > all_b) { all_b = nb; rall_b = r; hall_b = hb; }
    }
    // ---- which maxima carry the incoming gradient (model.py:82,100: the mask zeroes off rows) ----
    const float* gv = dvw + v * 128;
    float g_a = gv[lane], g_b = gv[lane + 32], gp_a = gv[64 + lane], gp_b = gv[96 + lane];
    if (!(any_on && (!any_off || on_a > 0.0f))) g_a = 0.0f;
    if (!(any_on && (!any_off || on_b > 0.0f))) g_b = 0.0f;
    if (!(any_on && (!any_off || all_a > 0.0f))) gp_a = 0.0f;
    if (!(any_on && (!any_off || all_b > 0.0f))) gp_b = 0.0f;
    if (PASS == 1) {
      dbe2a += g_a + gp_a;
      dbe2b += g_b + gp_b;
      dg2a = fmaf(g_a, (hon_a - m2a) * i2a, fmaf(gp_a, (hall_a - m2a) * i2a, dg2a));
      dg2b = fmaf(g_b, (hon_b - m2b) * i2b, fmaf(gp_b, (hall_b - m2b) * i2b, dg2b));
      __syncwarp();
      continue;
    }
    // ---- sweep 2: dz2 per row, dW2 / db2, gradient of the per-row half of the VFE-2 input ----
    float dzon_a = 0.0f, dzon_b = 0.0f;  // sum of dz2 over the rows with mask 1 (the agg1 half sees only this sum)
    for (int r = 0; r < R; ++r) {
      const bool on = (on_bits >> r) & 1;
      const float m = (r == nu) ? tail_mult : 1.0f;
      float pa = b2a, pb = b2b;
      const float* xr = n1 + r * 16;
      if (on) {
        pa = ca; pb = cb;
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          pa = fmaf(wa[k], xr[k], pa);
          pb = fmaf(wb[k], xr[k], pb);
        }
      }
      const float ha = fmaxf(pa, 0.0f), hb = fmaxf(pb, 0.0f);
      const float dna = (r == ron_a ? g_a : 0.0f) + (r == rall_a ? gp_a : 0.0f);
      const float dnb = (r == ron_b ? g_b : 0.0f) + (r == rall_b ? gp_b : 0.0f);
      const float dza = ha > 0.0f ? s2a * (dna - m * fmaf((ha - m2a) * i2a, B2a, A2a)) : 0.0f;
      const float dzb = hb > 0.0f ? s2b * (dnb - m * fmaf((hb - m2b) * i2b, B2b, A2b)) : 0.0f;
      db2a += dza;
      db2b += dzb;
      if (on) {  // warp-uniform
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          dWa[k] = fmaf(dza, xr[k], dWa[k]);
          dWb[k] = fmaf(dzb, xr[k], dWb[k]);
        }
        dzon_a += dza;
        dzon_b += dzb;
        dzs[lane] = dza;
        dzs[lane + 33] = dzb;
        __syncwarp();
        // d c1[r][k] for k < 16: lane = (k, half of the 64 units)
        float dc = 0.0f;
        const float* wk = S.w2lo + half * (32 * 16 + 16) + u1;
        const float* dzh = dzs + half * 33;
#pragma unroll 8
        for (int u = 0; u < 32; ++u) dc = fmaf(wk[u * 16], dzh[u], dc);
        dc += __shfl_xor_sync(0xffffffffu, dc, 16);
        if (lane < 16) g1[r * 16 + lane] = dc;
        __syncwarp();
      }
    }
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      dWa[16 + k] = fmaf(dzon_a, agg1[k], dWa[16 + k]);
      dWb[16 + k] = fmaf(dzon_b, agg1[k], dWb[16 + k]);
    }
    {  // agg1's gradient = W2[:,16:32]^T (sum of dz2 over the on rows); it goes to the row that attains the max
      dzs[lane] = dzon_a;
      dzs[lane + 33] = dzon_b;
      __syncwarp();
      float dc = 0.0f;
      const float* wk = S.w2hi + half * (32 * 16 + 16) + u1;
      const float* dzh = dzs + half * 33;
#pragma unroll 8
      for (int u = 0; u < 32; ++u) dc = fmaf(wk[u * 16], dzh[u], dc);
      dc += __shfl_xor_sync(0xffffffffu, dc, 16);
      if (lane < 16) g1[arg1 * 16 + lane] += dc;
    }
    __syncwarp();
    // ---- BatchNorm-1 sums; row gradients kept for PASS 3 ----
    for (int e = lane; e < R * 16; e += 32) {
      const float g = g1[e], h = h1s[e];
      dbe1 += g;
      dg1 = fmaf(g, (h - m1u) * i1u, dg1);
      grow[e] = g;
    }
    __syncwarp();
  }

  // ---- block reduction (shared fp32 atomics), then one fp64 atomic per value and block ----
  if (PASS == 1) {
    atomicAdd(&S.blk[kOffBe2 + lane], dbe2a);
    atomicAdd(&S.blk[kOffBe2 + lane + 32], dbe2b);
    atomicAdd(&S.blk[kOffG2 + lane], dg2a);
    atomicAdd(&S.blk[kOffG2 + lane + 32], dg2b);
  } else if (PASS == 2) {
#pragma unroll
    for (int k = 0; k < 32; ++k) {
      atomicAdd(&S.blk[kOffW2 + lane * 32 + k], dWa[k]);
      atomicAdd(&S.blk[kOffW2 + (lane + 32) * 32 + k], dWb[k]);
    }
    atomicAdd(&S.blk[kOffB2 + lane], db2a);
    atomicAdd(&S.blk[kOffB2 + lane + 32], db2b);
    atomicAdd(&S.blk[kOffBe1 + u1], dbe1);
    atomicAdd(&S.blk[kOffG1 + u1], dg1);
  } else {
#pragma unroll
    for (int j = 0; j < 7; ++j) atomicAdd(&S.blk[kOffW1 + u1 * 7 + j], dW1[j]);
    atomicAdd(&S.blk[kOffB1 + u1], db1);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kAccN; i += blockDim.x) {
    const float x = S.blk[i];
    if (x != 0.0f) atomicAdd(&acc_out[i], (double)x);
  }
}

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

int vn_bn_relu_bwd(const void* raw_bf16, const void* dy, int32_t dy_mode, int32_t B, int32_t D, int32_t H, int32_t W,
                   const float* prm, int32_t apply, double* sums, void* draw_bf16, vn_stream_t stream) {
  VN_REQUIRE(raw_bf16 && dy && prm, "bn_relu_bwd: null pointer");
  VN_REQUIRE(B >= 1 && D >= 1 && H >= 1 && W >= 1, "bn_relu_bwd: empty tensor");
  VN_REQUIRE(dy_mode == 0 || dy_mode == 1, "bn_relu_bwd: dy_mode=%d", dy_mode);
  VN_REQUIRE(apply ? draw_bf16 != nullptr : sums != nullptr, "bn_relu_bwd: missing output");
  VN_REQUIRE(dy_mode == 0 || W % 4 == 0, "bn_relu_bwd: W=%d must be a multiple of 4 for the (B, c*D+d, H, W) layout", W);
  VN_REQUIRE((reinterpret_cast<uintptr_t>(raw_bf16) & 15) == 0 && (reinterpret_cast<uintptr_t>(dy) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(draw_bf16) & 15) == 0,
             "bn_relu_bwd: buffers must be 16-byte aligned");
  cudaStream_t st = as_stream(stream);
  const int64_t P = (int64_t)B * D * H * W;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (dy_mode == 0) {
    const int64_t n_vec = P * 8;
    const int64_t want = (n_vec + 255) / 256;
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    if (apply)
      k_bn_bwd_ndhwc<true><<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(raw_bf16), reinterpret_cast<const uint4*>(dy),
                                                 n_vec, prm, nullptr, reinterpret_cast<uint4*>(draw_bf16));
    else
      k_bn_bwd_ndhwc<false><<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(raw_bf16), reinterpret_cast<const uint4*>(dy),
                                                  n_vec, prm, sums, nullptr);
  } else {
    const int64_t want = (int64_t)B * D * H * ((W + kFinW - 1) / kFinW);
    const unsigned grid = (unsigned)(want < cap ? want : cap);
    if (apply)
      k_bn_bwd_final<true><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(raw_bf16),
                                                 reinterpret_cast<const float*>(dy), B, D, H, W, prm, nullptr,
                                                 reinterpret_cast<__nv_bfloat16*>(draw_bf16));
    else
      k_bn_bwd_final<false><<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(raw_bf16),
                                                  reinterpret_cast<const float*>(dy), B, D, H, W, prm, sums, nullptr);
  }
  VN_LAUNCH_CHECK("k_bn_bwd");
  return VN_OK;
}

int vn_conv1_dgrad_gather(const void* draw_bf16, const void* weight_taps, const int64_t* coords, int64_t K, int32_t B,
                          int32_t Dr, int32_t H, int32_t W, int32_t stride_d, int32_t pad_d, float* dvw,
                          vn_stream_t stream) {
  VN_REQUIRE(K >= 0 && B >= 1 && Dr >= 1 && H >= 1 && W >= 1, "conv1_dgrad_gather: bad shape");
  VN_REQUIRE(stride_d == 1 || stride_d == 2, "conv1_dgrad_gather: stride_d=%d", stride_d);
  VN_REQUIRE(pad_d == 0 || pad_d == 1, "conv1_dgrad_gather: pad_d=%d", pad_d);
  if (K == 0) return VN_OK;
  VN_REQUIRE(draw_bf16 && weight_taps && coords && dvw, "conv1_dgrad_gather: null pointer");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(draw_bf16) & 15) == 0 && (reinterpret_cast<uintptr_t>(weight_taps) & 15) == 0,
             "conv1_dgrad_gather: buffers must be 16-byte aligned");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(dvw) & 15) == 0, "conv1_dgrad_gather: dvw must be 16-byte aligned");
  const int64_t blocks = (K + kGV - 1) / kGV;
  VN_REQUIRE(blocks < (int64_t)0x7fffffff, "conv1_dgrad_gather: too many voxels");
  VN_CUDA(cudaFuncSetAttribute(k_conv1_dgrad_gather, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGatherSmem));
  k_conv1_dgrad_gather<<<(unsigned)blocks, kGV, kGatherSmem, as_stream(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(draw_bf16), reinterpret_cast<const __nv_bfloat16*>(weight_taps),
      reinterpret_cast<const long long*>(coords), K, Dr, H, W, stride_d, pad_d, dvw);
  VN_LAUNCH_CHECK("k_conv1_dgrad_gather");
  return VN_OK;
}

int vn_svfe_bwd(const float* feature, int64_t K, int32_t T, const float* dvw, int32_t pass, const float* w1,
                const float* b1, const float* w2, const float* b2, const float* bn1, const float* bn2,
                float* row_scratch, double* acc, vn_stream_t stream) {
  VN_REQUIRE(K >= 0, "svfe_bwd: K < 0");
  VN_REQUIRE(T >= 1 && T <= 64, "svfe_bwd: T=%d outside [1,64]", T);
  VN_REQUIRE(pass >= 1 && pass <= 3, "svfe_bwd: pass=%d", pass);
  if (K == 0) return VN_OK;
  VN_REQUIRE(feature && dvw && w1 && b1 && w2 && b2 && bn1 && bn2 && acc, "svfe_bwd: null pointer");
  VN_REQUIRE(pass == 1 || row_scratch != nullptr, "svfe_bwd: passes 2 and 3 need row_scratch [K][T][16]");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(acc) & 7) == 0, "svfe_bwd: acc must be 8-byte aligned");
  cudaStream_t st = as_stream(stream);
  const int in_ld = (T * 7 + 3) & ~3;
  const size_t smem = sizeof(float) * (size_t)kBwWarps * (in_ld + (T + 1) * 16 + T * 16 + T * 16 + 72);
  const int64_t want = (K + kBwWarps - 1) / kBwWarps;
  const int64_t cap = (int64_t)sm_count() * 2;
  const unsigned grid = (unsigned)(want < cap ? want : cap);
#define VN_SVFE_BWD(P)                                                                                              \
  do {                                                                                                              \
    VN_CUDA(cudaFuncSetAttribute(k_svfe_bwd<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));           \
    k_svfe_bwd<P><<<grid, kBwWarps * 32, smem, st>>>(feature, K, T, dvw, w1, b1, w2, b2, bn1, bn2, row_scratch, acc);            \
  } while (0)
  if (pass == 1) VN_SVFE_BWD(1);
  else if (pass == 2) VN_SVFE_BWD(2);
  else VN_SVFE_BWD(3);
#undef VN_SVFE_BWD
  VN_LAUNCH_CHECK("k_svfe_bwd");
  return VN_OK;
}

}  // extern "C"
