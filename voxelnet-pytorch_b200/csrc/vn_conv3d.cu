// Conv3d(k=3) + BatchNorm3d(eval) + ReLU as a bf16 implicit GEMM on tcgen05 / TMEM, TMA-fed.
// Replaces ConvMD(3, ...).forward (reference voxelnet/model.py:158-167) for the three middle layers
// instantiated at model.py:207-209, and (out_mode 1) the reshape at model.py:262.
//
// GEMM view: M = output positions, N = 64 output channels, K = 27 taps x Cin.
//
// Tiling (DESIGN.md "Conv3d"): one CTA owns a 16 (h) x 8 (w) spatial tile for ALL output depths of
// one frame: Dout accumulators of 128 x 64 fp32 live side by side in TMEM (64 columns each, depth
// ascending).
//
// K loop: for each (64-channel chunk, kw) - 6 or 3 passes - the 72 KB of weights of that pass are
// made resident in shared memory (one bulk copy, double buffered) and the INPUT planes are walked in
// ascending order.  An input plane p feeds output depth d' = (p + pad_d - kd) / stride_d for every
// valid kd; those d' are consecutive, so the depth taps of one plane are fused along N: ONE
// tcgen05.mma of shape 128 x (64*nkd) x 16 multiplies the A tile with the nkd weight tiles laid side
// by side and lands in nkd adjacent accumulators.  This reads the A operand from shared memory once
// instead of nkd times (N = 64 alone is shared-memory-bandwidth bound: 6 KB of operands per 32
// tensor cycles).  Per tile the kernel pulls 432/216 KB of weights (each exactly once) plus one A
// box per (plane, chunk, kw) from L2.  In the last pass accumulators complete in ascending depth
// order, so the epilogue drains d' while the tensor core finishes d'+1.. and starts the next tile.
//
//   A operand: per (plane, 64-channel chunk, kw) ONE TMA box of 64 ch x 8 w x 18 h (an h-halo'd
//              tile, 128-byte rows, SWIZZLE_128B).  The three kh taps read it at +kh*1024 B: one
//              image row of the tile is exactly one 8-row swizzle atom, so every tap start stays
//              atom-aligned.  Out-of-range h / w / (skipped) d are the convolution's zero padding,
//              supplied by TMA's out-of-bounds zero fill.
//   B operand: weights pre-packed as [chunk][kw][kd class][kh][kd descending = d' ascending] K-major
//              [64 n x 64 k] tiles already in the swizzled shared-memory image.
//   Warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer (both run warp-uniform
//              control flow and elect one lane for the asynchronous instructions), warps 2..5 =
//              epilogue (TMEM -> registers -> scale/shift/ReLU -> global).
#include <cuda.h>
#include <cuda_bf16.h>

#include "vn_common.cuh"

namespace vn {
namespace {

constexpr int kTileH = 16, kTileW = 8;             // 128 output positions per accumulator
constexpr int kHaloH = kTileH + 2;
constexpr uint32_t kABytes = kHaloH * kTileW * 128;  // 18432: 144 rows of 64 bf16
constexpr uint32_t kBTapBytes = 64 * 128;            // 8192: [64 n][64 k] bf16
constexpr uint32_t kBSlotBytes = 9 * kBTapBytes;     // all weights of one (chunk, kw): 3 kh x 3 kd tiles = 72 KB
constexpr int kNA = 4, kNB = 2;                      // ring depths
constexpr int kMaxDout = 8;                          // 8 x 64 = 512 TMEM columns
constexpr int kMaxPlanes = 16;
constexpr int kThreads = 6 * 32;
constexpr uint32_t kSmemData = kNA * kABytes + kNB * kBSlotBytes;  // 221184
constexpr uint32_t kSmemBytes = kSmemData + 1024 /*align slack*/ + 1024 /*barriers, params*/;

// What one input plane contributes (host-computed; see plan_planes()).
struct PlaneInfo {
  int32_t nkd;        // number of valid depth taps (0: plane unused)
  int32_t dmin;       // first output depth fed; tap j feeds dmin + j
  int32_t new_acc;    // output depth whose FIRST contribution comes from this plane, or -1
  uint32_t woff;      // byte offset of this plane's first weight tile inside a (chunk, kw) weight group
  uint32_t wstride;   // bytes between the kh = 0, 1, 2 tile rows of the plane's kd class
  int32_t done_mask;  // bit d set: accumulator d is complete after this plane
};

struct ConvParams {
  int B, Din, Dout, H, W;
  int nchunk;      // Cin / 64
  int tiles_h, tiles_w, n_tiles;
  int out_mode;    // 0: bf16 NDHWC, 1: f32 [B, 64*Dout, H, W] with channel = c*Dout + d
  const uint8_t* wpack;
  const uint8_t* box_flags;  // [B][Din][tiles_h][tiles_w][3] or nullptr
  const float* scale;
  const float* shift;
  void* out;
  PlaneInfo plane[kMaxPlanes];
};

struct Barriers {
  uint64_t full_a[kNA], empty_a[kNA];
  uint64_t full_b[kNB], empty_b[kNB];
  uint64_t acc_full[kMaxDout], acc_empty[kMaxDout];
  uint32_t tmem_base;
  uint32_t pad;
  float scale[64], shift[64];
};
static_assert(sizeof(Barriers) <= 1024, "barrier block too large");

// Which input boxes of a tile can hold non-zero data: bit (p*3 + kw).  Without a flag table every
// box is live.  With one (written by vn_scatter_dense_bf16 from the voxel coordinates), a zero flag
// guarantees the 18 x 8 box of plane p at w-shift kw is all zeros, so its TMA load and MMAs are
// skipped - exact, since zero operands contribute nothing.  All 32 lanes must call this.
template <bool kSkip>
__device__ __forceinline__ uint64_t live_boxes(const ConvParams& P, int b, int th, int tw, uint32_t lane) {
  if (!kSkip) return ~0ull;
  uint64_t m = 0;
  if ((int)lane < P.Din) {
    const uint8_t* f = P.box_flags + ((((size_t)b * P.Din + lane) * P.tiles_h + th) * P.tiles_w + tw) * 3;
    const uint32_t bits = (f[0] ? 1u : 0u) | (f[1] ? 2u : 0u) | (f[2] ? 4u : 0u);
    m = (uint64_t)bits << (3 * lane);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) m |= __shfl_xor_sync(0xffffffffu, m, d);
  return m;
}

// Accumulators (bit d) that receive at least one live box in this tile.
__device__ __forceinline__ uint32_t touched_accumulators(const ConvParams& P, uint64_t live) {
  uint32_t t = 0;
  for (int p = 0; p < P.Din; ++p) {
    const PlaneInfo& pi = P.plane[p];
    if (pi.nkd != 0 && ((live >> (p * 3)) & 7)) t |= ((1u << pi.nkd) - 1u) << pi.dmin;
  }
  return t;
}

// kSkip = false compiles the occupancy logic away for dense inputs (conv-2, conv-3, drop-in calls).
template <bool kSkip>
__global__ void __launch_bounds__(kThreads, 1)
    k_conv3d_bn_relu(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ ConvParams P) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B operands need 1024-byte aligned tiles
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kNA * kABytes;
  Barriers* bar = reinterpret_cast<Barriers*>(smem + kSmemData);

  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t acc_cols = (uint32_t)P.Dout * 64;
  const uint32_t tmem_cols = acc_cols <= 32 ? 32 : (acc_cols <= 64 ? 64 : (acc_cols <= 128 ? 128 : (acc_cols <= 256 ? 256 : 512)));

  if (threadIdx.x == 0) {
    for (int i = 0; i < kNA; ++i) { mbar_init(&bar->full_a[i], 1); mbar_init(&bar->empty_a[i], 1); }
    for (int i = 0; i < kNB; ++i) { mbar_init(&bar->full_b[i], 1); mbar_init(&bar->empty_b[i], 1); }
    for (int i = 0; i < kMaxDout; ++i) { mbar_init(&bar->acc_full[i], 1); mbar_init(&bar->acc_empty[i], 4); }
    mbar_fence_init();
    fence_proxy_async_smem();
    tma_prefetch_desc(&tmap_x);
  }
  if (threadIdx.x >= 64 && threadIdx.x < 128) {
    bar->scale[threadIdx.x - 64] = P.scale[threadIdx.x - 64];
    bar->shift[threadIdx.x - 64] = P.shift[threadIdx.x - 64];
  }
  if (warp == 1) {
    tmem_alloc(&bar->tmem_base, tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = bar->tmem_base;

  if (warp == 0) {
    // ===================== TMA producer (warp-uniform; one elected lane issues) =====================
    uint32_t a_it = 0, b_it = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
      const int tw = tile % P.tiles_w;
      const int th = (tile / P.tiles_w) % P.tiles_h;
      const int b = tile / (P.tiles_w * P.tiles_h);
      const int h0 = th * kTileH, w0 = tw * kTileW;
      const uint64_t live = live_boxes<kSkip>(P, b, th, tw, lane);
      for (int c = 0; c < P.nchunk; ++c) {
        for (int kw = 0; kw < 3; ++kw) {
          // all weights of this (chunk, kw): resident while the K loop walks the input planes
          const uint32_t sb = b_it % kNB;
          mbar_wait(&bar->empty_b[sb], ((b_it / kNB) & 1) ^ 1);
          if (elect_one()) {
            mbar_arrive_expect_tx(&bar->full_b[sb], kBSlotBytes);
            bulk_load_1d(smem_b + sb * kBSlotBytes, P.wpack + (size_t)(c * 3 + kw) * kBSlotBytes, kBSlotBytes,
                         &bar->full_b[sb]);
          }
          __syncwarp();
          ++b_it;
          for (int p = 0; p < P.Din; ++p) {
            if (P.plane[p].nkd == 0 || !((live >> (p * 3 + kw)) & 1)) continue;
            const uint32_t sa = a_it % kNA;
            mbar_wait(&bar->empty_a[sa], ((a_it / kNA) & 1) ^ 1);
            if (elect_one()) {
              mbar_arrive_expect_tx(&bar->full_a[sa], kABytes);
              tma_load_5d(smem_a + sa * kABytes, &tmap_x, &bar->full_a[sa], c * 64, w0 + kw - 1, h0 - 1, p, b);
            }
            __syncwarp();
            ++a_it;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (warp-uniform; one elected lane issues) =====================
    uint32_t a_it = 0, b_it = 0;
    uint32_t acc_phase = 0;  // bit d: parity of the number of tiles that used accumulator d so far
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
      const int tw = tile % P.tiles_w;
      const int th = (tile / P.tiles_w) % P.tiles_h;
      const int b = tile / (P.tiles_w * P.tiles_h);
      const uint64_t live = live_boxes<kSkip>(P, b, th, tw, lane);
      const uint32_t touched = touched_accumulators(P, live);
      uint32_t started = 0;  // accumulators that already hold a partial sum of this tile
      for (int c = 0; c < P.nchunk; ++c) {
        for (int kw = 0; kw < 3; ++kw) {
          const bool last_pass = (c == P.nchunk - 1) & (kw == 2);
          const uint32_t sb = b_it % kNB;
          mbar_wait(&bar->full_b[sb], (b_it / kNB) & 1);
          tc_fence_after_sync();
          const uint32_t b_addr = smem_u32(smem_b + sb * kBSlotBytes);
          for (int p = 0; p < P.Din; ++p) {
            const PlaneInfo& pi = P.plane[p];
            if (pi.nkd == 0) continue;
            if ((live >> (p * 3 + kw)) & 1) {
              const uint32_t range = ((1u << pi.nkd) - 1u) << pi.dmin;
              const uint32_t fresh = range & ~started;
              // Accumulators first touched here must have been drained by the epilogue, and need
              // accumulate=0 while already-started neighbours need 1: the first k-step is then issued
              // per accumulator (N = 64) instead of fused along N.
              if (fresh) {
                for (int d = pi.dmin; d < pi.dmin + pi.nkd; ++d)
                  if ((fresh >> d) & 1) mbar_wait(&bar->acc_empty[d], ((acc_phase >> d) & 1) ^ 1);
                tc_fence_after_sync();
              }
              const uint32_t d_tmem = tmem_base + (uint32_t)pi.dmin * 64;
              const uint32_t idesc_all = umma_idesc_bf16(128, 64 * pi.nkd);
              const uint32_t sa = a_it % kNA;
              mbar_wait(&bar->full_a[sa], (a_it / kNA) & 1);
              tc_fence_after_sync();
              const uint32_t a_addr = smem_u32(smem_a + sa * kABytes);
              if (elect_one()) {
#pragma unroll
                for (int kh = 0; kh < 3; ++kh) {
                  const uint64_t da0 = umma_desc_sw128(a_addr + kh * (kTileW * 128), 1024);
                  const uint64_t db0 = umma_desc_sw128(b_addr + pi.woff + kh * pi.wstride, 1024);
                  int k0 = 0;
                  if (kh == 0 && fresh) {
                    for (int j = 0; j < pi.nkd; ++j)
                      umma_bf16_ss(d_tmem + j * 64, da0, db0 + (uint64_t)((j * kBTapBytes) >> 4), umma_idesc_bf16(128, 64),
                                   (started >> (pi.dmin + j)) & 1);
                    k0 = 1;
                  }
#pragma unroll
                  for (int k = 0; k < 4; ++k) {
                    if (k >= k0) umma_bf16_ss(d_tmem, da0 + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), idesc_all, 1);
                  }
                }
                umma_commit(&bar->empty_a[sa]);
              }
              __syncwarp();
              started |= range;
              ++a_it;
            }
            // accumulators whose last possible contribution is this plane of the last pass
            if (last_pass && (pi.done_mask & touched) != 0) {
              if (elect_one()) {
                for (int d = 0; d < P.Dout; ++d)
                  if (((pi.done_mask & touched) >> d) & 1) umma_commit(&bar->acc_full[d]);
              }
              __syncwarp();
            }
          }
          if (elect_one()) umma_commit(&bar->empty_b[sb]);
          __syncwarp();
          ++b_it;
        }
      }
      acc_phase ^= touched;
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const uint32_t quad = warp & 3;                // TMEM lane quadrant this warp may read
    const uint32_t m = quad * 32 + lane;           // accumulator row = output position in the tile
    const int hl = m / kTileW, wl = m % kTileW;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x) {
      const int tw = tile % P.tiles_w;
      const int th = (tile / P.tiles_w) % P.tiles_h;
      const int b = tile / (P.tiles_w * P.tiles_h);
      const int h = th * kTileH + hl, w = tw * kTileW + wl;
      const bool inside = h < P.H && w < P.W;
      const uint32_t touched = touched_accumulators(P, live_boxes<kSkip>(P, b, th, tw, lane));
      for (int dout = 0; dout < P.Dout; ++dout) {
        const bool has_acc = (touched >> dout) & 1;  // else: no non-zero input reached this accumulator
        if (has_acc) {
          mbar_wait(&bar->acc_full[dout], (acc_phase >> dout) & 1);
          tc_fence_after_sync();
        }
        const uint32_t taddr = tmem_base + ((quad * 32) << 16) + (uint32_t)dout * 64;
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t r[32];
          if (has_acc) {
            tmem_ld_32x32b_x32(taddr + half * 32, r);
            tmem_ld_wait();
            if (half == 1) {
              // accumulator fully read: hand it back to the MMA issuer before the global stores
              tc_fence_before_sync();
              __syncwarp();
              if (lane == 0) mbar_arrive(&bar->acc_empty[dout]);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0u;
          }
          if (inside) {
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int c = half * 32 + j;
              v[j] = fmaxf(fmaf(__uint_as_float(r[j]), bar->scale[c], bar->shift[c]), 0.0f);
            }
            if (P.out_mode == 0) {
              __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(P.out) +
                                   ((((size_t)b * P.Dout + dout) * P.H + h) * P.W + w) * 64 + half * 32;
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                uint32_t pk[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const __nv_bfloat162 t2 = __floats2bfloat162_rn(v[q * 8 + 2 * e], v[q * 8 + 2 * e + 1]);
                  pk[e] = *reinterpret_cast<const uint32_t*>(&t2);
                }
                *reinterpret_cast<uint4*>(dst + q * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
              }
            } else {
              const size_t plane = (size_t)P.H * P.W;
              float* dst = reinterpret_cast<float*>(P.out) +
                           ((size_t)b * 64 * P.Dout + (size_t)(half * 32) * P.Dout + dout) * plane + (size_t)h * P.W + w;
#pragma unroll
              for (int j = 0; j < 32; ++j) dst[(size_t)j * P.Dout * plane] = v[j];
            }
          }
        }
      }
      acc_phase ^= touched;
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, tmem_cols);
  }
}

// Weight image.  Depth taps are grouped into "kd classes" that can be valid for the same input plane:
//   stride_d 1: one class {2,1,0};  stride_d 2: class 0 = {2,0} (planes with p+pad even), class 1 = {1}.
// Layout: [chunk][kw][class][kh][j: kd descending][n = 64][k = 64] bf16, rows 128B-swizzled: the 72 KB
// of one (chunk, kw) are contiguous (one bulk copy), and the taps valid for a plane sit side by side.
struct KdClasses {
  int n_classes;
  int nkd[2];
  int kd[2][3];
  uint32_t base[2];  // byte offset of the class inside a (chunk, kw) group
};

__host__ __device__ inline KdClasses kd_classes(int stride_d, int cin) {
  KdClasses k{};
  (void)cin;
  if (stride_d == 1) {
    k.n_classes = 1;
    k.nkd[0] = 3; k.kd[0][0] = 2; k.kd[0][1] = 1; k.kd[0][2] = 0;
    k.base[0] = 0;
  } else {
    k.n_classes = 2;
    k.nkd[0] = 2; k.kd[0][0] = 2; k.kd[0][1] = 0;
    k.nkd[1] = 1; k.kd[1][0] = 1;
    k.base[0] = 0;
    k.base[1] = 3 * 2 * kBTapBytes;  // after class 0: 3 kh x 2 taps
  }
  return k;
}

// torch [64][cin][3][3][3] fp32 -> the image above.
__global__ void __launch_bounds__(256)
    k_pack_weights(const float* __restrict__ w, int cin, int stride_d, uint8_t* __restrict__ packed) {
  const int nchunk = cin / 64;
  const KdClasses kc = kd_classes(stride_d, cin);
  const int total = 27 * cin * 64;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    int r = i;
    const int k = r % 64; r /= 64;
    const int n = r % 64; r /= 64;
    const int kh = r % 3; r /= 3;
    const int kw = r % 3; r /= 3;
    const int c = r % nchunk; r /= nchunk;
    const int kd = r;
    int cls = 0, j = 0;
    for (int a = 0; a < kc.n_classes; ++a)
      for (int q = 0; q < kc.nkd[a]; ++q)
        if (kc.kd[a][q] == kd) { cls = a; j = q; }
    const float v = w[((((size_t)n * cin + (c * 64 + k)) * 3 + kd) * 3 + kh) * 3 + kw];
    const size_t tile = (size_t)(c * 3 + kw) * kBSlotBytes + kc.base[cls] + (size_t)(kh * kc.nkd[cls] + j) * kBTapBytes;
    const size_t off = tile + (size_t)n * 128 + ((((k >> 3) ^ (n & 7)) << 4) | ((k & 7) << 1));
    *reinterpret_cast<__nv_bfloat16*>(packed + off) = __float2bfloat16_rn(v);
  }
}

// Per-plane schedule: which taps are valid, where their weights sit, which accumulators start/finish.
int plan_planes(ConvParams& P, int cin, int stride_d, int pad_d) {
  const KdClasses kc = kd_classes(stride_d, cin);
  int first_plane[kMaxDout], last_plane[kMaxDout];
  for (int d = 0; d < P.Dout; ++d) { first_plane[d] = -1; last_plane[d] = -1; }
  for (int p = 0; p < P.Din; ++p) {
    PlaneInfo& pi = P.plane[p];
    pi = PlaneInfo{0, 0, -1, 0, 0, 0};
    const int cls = stride_d == 1 ? 0 : (((p + pad_d) % 2 == 0) ? 0 : 1);
    int jfirst = -1, jlast = -1, dmin = 0;
    for (int j = 0; j < kc.nkd[cls]; ++j) {  // kd descending -> d ascending
      const int num = p + pad_d - kc.kd[cls][j];
      if (num < 0 || num % stride_d != 0) continue;
      const int d = num / stride_d;
      if (d >= P.Dout) continue;
      if (jfirst < 0) { jfirst = j; dmin = d; }
      if (jlast >= 0 && j != jlast + 1) return -1;  // taps must be contiguous
      jlast = j;
      if (d != dmin + (j - jfirst)) return -1;      // and feed consecutive depths
      if (first_plane[d] < 0) first_plane[d] = p;
      last_plane[d] = p;
    }
    if (jfirst < 0) continue;
    pi.nkd = jlast - jfirst + 1;
    pi.dmin = dmin;
    pi.woff = kc.base[cls] + (uint32_t)jfirst * kBTapBytes;
    pi.wstride = (uint32_t)kc.nkd[cls] * kBTapBytes;
  }
  for (int d = 0; d < P.Dout; ++d) {
    if (first_plane[d] < 0) return -1;
    PlaneInfo& pf = P.plane[first_plane[d]];
    if (pf.new_acc >= 0) return -1;                      // at most one new accumulator per plane
    if (d != pf.dmin + pf.nkd - 1) return -1;            // and it is the plane's last tap
    pf.new_acc = d;
    P.plane[last_plane[d]].done_mask |= 1 << d;
  }
  return 0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

size_t vn_conv3d_box_flags_bytes(int32_t B, int32_t Din, int32_t H, int32_t W) {
  if (B < 1 || Din < 1 || H < 1 || W < 1 || W % kTileW != 0) return 0;
  return (size_t)B * Din * ((H + kTileH - 1) / kTileH) * (W / kTileW) * 3;
}

size_t vn_conv3d_packed_weight_bytes(int32_t cin) {
  if (cin != 64 && cin != 128) return 0;
  return (size_t)27 * cin * 64 * 2;
}

int vn_conv3d_pack_weights(const float* weight, int32_t cin, int32_t stride_d, void* packed, vn_stream_t stream) {
  VN_REQUIRE(cin == 64 || cin == 128, "conv3d_pack_weights: cin=%d (supported: 64, 128)", cin);
  VN_REQUIRE(stride_d == 1 || stride_d == 2, "conv3d_pack_weights: stride_d=%d (supported: 1, 2)", stride_d);
  VN_REQUIRE(weight != nullptr && packed != nullptr, "conv3d_pack_weights: null pointer");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0, "conv3d_pack_weights: packed must be 16-byte aligned");
  k_pack_weights<<<sm_count() * 4, 256, 0, as_stream(stream)>>>(weight, cin, stride_d, reinterpret_cast<uint8_t*>(packed));
  VN_LAUNCH_CHECK("k_pack_weights");
  return VN_OK;
}

int vn_conv3d_bn_relu_fwd(const void* x, int32_t B, int32_t Din, int32_t H, int32_t W, int32_t cin, int32_t stride_d,
                          int32_t pad_d, const void* packed_weight, const float* scale, const float* shift,
                          int32_t out_mode, void* y, const uint8_t* box_flags, vn_stream_t stream) {
  VN_REQUIRE(x && packed_weight && scale && shift && y, "conv3d: null pointer");
  VN_REQUIRE(cin == 64 || cin == 128, "conv3d: cin=%d (supported: 64, 128)", cin);
  VN_REQUIRE(B >= 1 && Din >= 1 && H >= 1 && W >= 1, "conv3d: empty input");
  VN_REQUIRE((stride_d == 1 || stride_d == 2) && (pad_d == 0 || pad_d == 1), "conv3d: stride_d=%d pad_d=%d unsupported",
             stride_d, pad_d);
  VN_REQUIRE(out_mode == 0 || out_mode == 1, "conv3d: out_mode=%d", out_mode);
  VN_REQUIRE(W % kTileW == 0, "conv3d: W=%d must be a multiple of %d", W, kTileW);
  VN_REQUIRE(Din <= kMaxPlanes, "conv3d: Din=%d exceeds %d", Din, kMaxPlanes);
  const int Dout = (Din + 2 * pad_d - 3) / stride_d + 1;
  VN_REQUIRE(Din + 2 * pad_d >= 3 && Dout >= 1 && Dout <= kMaxDout, "conv3d: Dout=%d outside [1,%d]", Dout, kMaxDout);
  VN_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(packed_weight) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(y) & 15) == 0,
             "conv3d: buffers must be 16-byte aligned");

  EncodeTiledFn encode = encode_tiled_fn();
  if (encode == nullptr) {
    set_error("conv3d: cuTensorMapEncodeTiled not available from the driver");
    return VN_ERR_DRIVER;
  }
  CUtensorMap tmap;
  const cuuint64_t gdim[5] = {(cuuint64_t)cin, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Din, (cuuint64_t)B};
  const cuuint64_t gstr[4] = {(cuuint64_t)cin * 2, (cuuint64_t)W * cin * 2, (cuuint64_t)H * W * cin * 2,
                              (cuuint64_t)Din * H * W * cin * 2};
  const cuuint32_t box[5] = {64, (cuuint32_t)kTileW, (cuuint32_t)kHaloH, 1, 1};
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  const CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(x), gdim, gstr, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) {
    set_error("conv3d: cuTensorMapEncodeTiled failed with CUresult %d", (int)cr);
    return VN_ERR_DRIVER;
  }

  ConvParams P{};
  P.B = B; P.Din = Din; P.Dout = Dout; P.H = H; P.W = W;
  P.nchunk = cin / 64;
  P.tiles_h = (H + kTileH - 1) / kTileH;
  P.tiles_w = W / kTileW;
  P.n_tiles = B * P.tiles_h * P.tiles_w;
  P.out_mode = out_mode;
  P.wpack = reinterpret_cast<const uint8_t*>(packed_weight);
  P.box_flags = box_flags;
  P.scale = scale; P.shift = shift;
  P.out = y;
  if (plan_planes(P, cin, stride_d, pad_d) != 0) {
    set_error("conv3d: unsupported depth schedule (Din=%d stride_d=%d pad_d=%d)", Din, stride_d, pad_d);
    return VN_ERR_UNSUPPORTED;
  }

  const int grid = P.n_tiles < sm_count() ? P.n_tiles : sm_count();
  if (box_flags != nullptr) {
    VN_CUDA(cudaFuncSetAttribute(k_conv3d_bn_relu<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    k_conv3d_bn_relu<true><<<grid, kThreads, kSmemBytes, as_stream(stream)>>>(tmap, P);
  } else {
    VN_CUDA(cudaFuncSetAttribute(k_conv3d_bn_relu<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    k_conv3d_bn_relu<false><<<grid, kThreads, kSmemBytes, as_stream(stream)>>>(tmap, P);
  }
  VN_LAUNCH_CHECK("k_conv3d_bn_relu");
  return VN_OK;
}

}  // extern "C"
