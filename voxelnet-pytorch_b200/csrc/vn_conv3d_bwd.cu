// Backward kernels of the Conv3d middle layers (training path; the reference gets them from autograd over
// nn.Conv3d / nn.BatchNorm3d / F.relu, voxelnet/model.py:158-167, driven by loss.backward() at train.py:149).
//
// k_conv3d_wgrad: dW[n][c][kd][kh][kw] = sum over (b, d', h, w) of dY[b,d',h,w,n] * X[b, d'*s-pad+kd, h+kh-1, w+kw-1, c]
// as a bf16 GEMM on tcgen05 with the POSITIONS as the K dimension.  Both operands sit in shared memory exactly
// as the forward kernel's TMA boxes (rows = positions, 64 channels = 128 bytes, SWIZZLE_128B) and are read as
// MN-major operands (channels are the M / N index):
//   A = X box, M = 128 = two 64-channel atoms `LBO` bytes apart.  The kh taps of one box are +1024 B apart (one
//       image row of the 16x8 tile = one swizzle atom), the kw taps / channel chunks are separate boxes, so a pair
//       of taps (cin = 64) or the two channel chunks of one tap (cin = 128) form one M = 128 operand by choosing LBO.
//   B = dY tile (no halo), N = 64 output channels.  K = 16 positions = two tile rows per tcgen05.mma.
// Accumulators [128 x 64] fp32 stay in TMEM over the CTA's whole slice of positions; a CTA owns one tap group
// (cin = 64: the 9 (kh,kw) taps of one kd = 5 accumulators; cin = 128: the 3 kh taps of one (kd,kw) = 3
// accumulators) and 1/S of the (frame, depth, tile) units.  At the end the partial sums are added to the global
// fp32 gradient with red.global.add.
#include <cuda.h>
#include <cuda_bf16.h>

#include "vn_common.cuh"

namespace vn {
namespace {

constexpr int kTileH = 16, kTileW = 8;
constexpr uint32_t kYBytes = kTileH * kTileW * 128;        // 16384: dY tile, 128 positions x 64 ch
constexpr uint32_t kXBytes = (kTileH + 2) * kTileW * 128;  // 18432: h-halo'd X box
constexpr uint32_t kStageBytes = kYBytes + 3 * kXBytes;    // 71680
constexpr int kStages = 3;
constexpr int kMaxOps = 5;
constexpr int kWgThreads = 6 * 32;  // producer, MMA issuer, 4 epilogue warps

struct WgParams {
  int B, Din, Dout, H, W, cin, stride_d, pad_d;
  int tiles_h, tiles_w;
  int n_groups, n_slices;
  float* dw;  // [27][cin][64] fp32, accumulated into
  // live-box table of the layer input ([B][Din][tiles_h][tiles_w][3], vn_conv3d_mark_boxes) or nullptr: a unit
  // whose X boxes hold only zeros contributes nothing and is skipped (exact for the voxel grid of layer 1)
  const uint8_t* flags;
};

struct WgControl {
  uint64_t full[kStages], empty[kStages];
  uint64_t done;
  uint32_t tmem_base;
  uint32_t pad;
};
constexpr uint32_t kWgSmem = kStages * kStageBytes + 1024 + (uint32_t)sizeof(WgControl);
static_assert(kWgSmem <= 232448, "shared memory budget");

// One tcgen05.mma per k-step: A = stage X area + a_off with atom stride lbo; accumulator `slot`;
// rows 0-63 / 64-127 of the accumulator are (tap[0], channel base cb[0]) / (tap[1], cb[1]); tap < 0: unused.
struct WgOp {
  uint32_t a_off, lbo;
  int tap[2], cb[2];
};
struct WgGroup {
  int kd, nboxes, nops;
  int box_c[3], box_kw[3];
  WgOp op[kMaxOps];
};

__device__ __forceinline__ int tap_index(int kd, int kh, int kw) { return (kd * 3 + kh) * 3 + kw; }

__device__ __forceinline__ WgGroup make_group(int cin, int g) {
  WgGroup G{};
  if (cin == 64) {  // g = kd; boxes kw = 0,1,2; accumulators: (kh0|kh1) per kw, (kh2,kw0 | kh2,kw1), (kh2,kw2 | -)
    G.kd = g;
    G.nboxes = 3;
    for (int i = 0; i < 3; ++i) { G.box_c[i] = 0; G.box_kw[i] = i; }
    G.nops = 5;
    for (int kw = 0; kw < 3; ++kw)
      G.op[kw] = WgOp{(uint32_t)kw * kXBytes, 1024u, {tap_index(g, 0, kw), tap_index(g, 1, kw)}, {0, 0}};
    G.op[3] = WgOp{2 * 1024u, kXBytes, {tap_index(g, 2, 0), tap_index(g, 2, 1)}, {0, 0}};
    G.op[4] = WgOp{2 * kXBytes + 2 * 1024u, 0u, {tap_index(g, 2, 2), -1}, {0, 0}};
  } else {  // cin == 128: g = kd*3 + kw; boxes = the two channel chunks; one accumulator per kh
    G.kd = g / 3;
    const int kw = g % 3;
    G.nboxes = 2;
    G.box_c[0] = 0; G.box_c[1] = 64;
    G.box_kw[0] = kw; G.box_kw[1] = kw;
    G.nops = 3;
    for (int kh = 0; kh < 3; ++kh)
      G.op[kh] = WgOp{(uint32_t)kh * 1024u, kXBytes, {tap_index(G.kd, kh, kw), tap_index(G.kd, kh, kw)}, {0, 64}};
  }
  return G;
}

// SW128 MN-major shared-memory descriptor: 64-element (128 B) atoms along M/N `lbo` bytes apart, 8-row groups
// along K `sbo` bytes apart (cute::UMMA canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units).
__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

__global__ void __launch_bounds__(kWgThreads, 1)
    k_conv3d_wgrad(const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_dy,
                   const __grid_constant__ WgParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  WgControl* ctl = reinterpret_cast<WgControl*>(smem + kStages * kStageBytes);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  const int g = blockIdx.x % P.n_groups, slice = blockIdx.x / P.n_groups;
  const WgGroup G = make_group(P.cin, g);
  const int64_t n_units = (int64_t)P.B * P.Dout * P.tiles_h * P.tiles_w;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&ctl->full[i], 1); mbar_init(&ctl->empty[i], 1); }
    mbar_init(&ctl->done, 1);
    mbar_fence_init();
    fence_proxy_async_smem();
    tma_prefetch_desc(&tmap_x);
    tma_prefetch_desc(&tmap_dy);
  }
  if (warp == 1) {
    tmem_alloc(&ctl->tmem_base, 512);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = ctl->tmem_base;

  // Units of this CTA: u = slice, slice + S, ...; unit -> (b, d', th, tw); X plane p = d'*s - pad + G.kd.
  auto geometry = [&](int64_t u, int& b, int& d, int& th, int& tw, int& p) {
    tw = (int)(u % P.tiles_w); u /= P.tiles_w;
    th = (int)(u % P.tiles_h); u /= P.tiles_h;
    d = (int)(u % P.Dout);
    b = (int)(u / P.Dout);
    p = d * P.stride_d - P.pad_d + G.kd;
  };
  // The warp tests 32 candidate units at once (one per lane: the live-box flags are a global-memory read each)
  // and returns the mask of those that have work; producer and issuer walk identical sequences.
  auto live_mask = [&](int64_t u0) -> uint32_t {
    const int64_t u = u0 + (int64_t)lane * P.n_slices;
    bool ok = false;
    if (u < n_units) {
      int b, d, th, tw, p;
      geometry(u, b, d, th, tw, p);
      ok = p >= 0 && p < P.Din;
      if (ok && P.flags != nullptr) {
        const uint8_t* f = P.flags + ((((size_t)b * P.Din + p) * P.tiles_h + th) * P.tiles_w + tw) * 3;
        ok = false;
        for (int i = 0; i < G.nboxes; ++i) ok |= f[G.box_kw[i]] != 0;
      }
    }
    return __ballot_sync(0xffffffffu, ok);
  };

  if (warp == 0) {
    // ================= TMA producer =================
    uint32_t it = 0;
    const uint32_t bytes = kYBytes + (uint32_t)G.nboxes * kXBytes;
    for (int64_t u0 = slice; u0 < n_units; u0 += 32 * (int64_t)P.n_slices)
    for (uint32_t todo = live_mask(u0); todo != 0; todo &= todo - 1) {
      int b, d, th, tw, p;
      geometry(u0 + (int64_t)(__ffs(todo) - 1) * P.n_slices, b, d, th, tw, p);
      const uint32_t s = it % kStages;
      mbar_wait(&ctl->empty[s], ((it / kStages) & 1) ^ 1);
      if (elect_one()) {
        uint8_t* st = smem + s * kStageBytes;
        mbar_arrive_expect_tx(&ctl->full[s], bytes);
        tma_load_5d(st, &tmap_dy, &ctl->full[s], 0, tw * kTileW, th * kTileH, d, b);
        for (int i = 0; i < G.nboxes; ++i)
          tma_load_5d(st + kYBytes + i * kXBytes, &tmap_x, &ctl->full[s], G.box_c[i], tw * kTileW + G.box_kw[i] - 1,
                      th * kTileH - 1, p, b);
      }
      __syncwarp();
      ++it;
    }
  } else if (warp == 1) {
    // ================= MMA issuer =================
    uint32_t it = 0;
    // M = 128, N = 64, bf16 x bf16 -> f32, A and B MN-major (bits 15, 16)
    const uint32_t idesc = umma_idesc_bf16(128, 64) | (1u << 15) | (1u << 16);
    for (int64_t u0 = slice; u0 < n_units; u0 += 32 * (int64_t)P.n_slices)
    for (uint32_t todo = live_mask(u0); todo != 0; todo &= todo - 1) {
      const uint32_t s = it % kStages;
      mbar_wait(&ctl->full[s], (it / kStages) & 1);
      tc_fence_after_sync();
      if (elect_one()) {
        const uint32_t y_addr = smem_u32(smem + s * kStageBytes);
        const uint32_t x_addr = y_addr + kYBytes;
#pragma unroll 1
        for (int ks = 0; ks < 8; ++ks) {  // 16 positions (two tile rows) per step
          const uint64_t bdesc = desc_mn_sw128(y_addr + ks * 2048, 0u, 1024u);
#pragma unroll
          for (int o = 0; o < kMaxOps; ++o) {
            if (o < G.nops) {
              const uint64_t adesc = desc_mn_sw128(x_addr + G.op[o].a_off + ks * 2048, G.op[o].lbo, 1024u);
              umma_bf16_ss(tmem_base + o * 64, adesc, bdesc, idesc, (it == 0 && ks == 0) ? 0u : 1u);
            }
          }
        }
        umma_commit(&ctl->empty[s]);
      }
      __syncwarp();
      ++it;
    }
    if (lane == 0) {
      ctl->pad = it;  // 0: nothing accumulated (read by the epilogue after the barrier)
      __threadfence_block();
    }
    __syncwarp();
    if (elect_one()) {
      if (it != 0) umma_commit(&ctl->done);
      else mbar_arrive(&ctl->done);
    }
    __syncwarp();
  } else {
    // ================= epilogue: TMEM -> global fp32 gradient =================
    mbar_wait(&ctl->done, 0);
    tc_fence_after_sync();
    const uint32_t n_done = *reinterpret_cast<volatile uint32_t*>(&ctl->pad);
    if (n_done != 0) {
      const uint32_t quad = warp & 3;              // TMEM lanes 32*quad .. +31
      const int row = quad * 32 + lane;            // accumulator row
      const int atom = row >> 6, ch = row & 63;
      for (int o = 0; o < G.nops; ++o) {
        const int tap = G.op[o].tap[atom];
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint32_t r[32];
          tmem_ld_32x32b_x32(tmem_base + ((quad * 32) << 16) + o * 64 + half * 32, r);
          tmem_ld_wait();
          if (tap >= 0) {
            float* dst = P.dw + ((size_t)tap * P.cin + G.op[o].cb[atom] + ch) * 64 + half * 32;
#pragma unroll
            for (int j = 0; j < 32; ++j) atomicAdd(dst + j, __uint_as_float(r[j]));
          }
        }
      }
    }
  }

  tc_fence_before_sync();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, 512);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
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

int encode_ndhwc(CUtensorMap* map, const void* base, int B, int D, int H, int W, int C, int box_h) {
  EncodeTiledFn enc = encode_fn();
  if (enc == nullptr) {
    set_error("conv3d backward: cuTensorMapEncodeTiled not available from the driver");
    return VN_ERR_DRIVER;
  }
  const cuuint64_t gdim[5] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)D, (cuuint64_t)B};
  const cuuint64_t gstr[4] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2, (cuuint64_t)D * H * W * C * 2};
  const cuuint32_t box[5] = {64, (cuuint32_t)kTileW, (cuuint32_t)box_h, 1, 1};
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  const CUresult cr = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), gdim, gstr, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (cr != CUDA_SUCCESS) {
    set_error("conv3d backward: cuTensorMapEncodeTiled failed with CUresult %d", (int)cr);
    return VN_ERR_DRIVER;
  }
  return VN_OK;
}

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

int vn_conv3d_wgrad(const void* x, const void* dy, int32_t B, int32_t Din, int32_t H, int32_t W, int32_t cin,
                    int32_t stride_d, int32_t pad_d, float* dw_taps, const uint8_t* box_flags, vn_stream_t stream) {
  VN_REQUIRE(x && dy && dw_taps, "conv3d_wgrad: null pointer");
  VN_REQUIRE(cin == 64 || cin == 128, "conv3d_wgrad: cin=%d (supported: 64, 128)", cin);
  VN_REQUIRE(B >= 1 && Din >= 1 && H >= 1 && W >= 1 && W % kTileW == 0, "conv3d_wgrad: bad shape");
  VN_REQUIRE((stride_d == 1 || stride_d == 2) && (pad_d == 0 || pad_d == 1) && Din + 2 * pad_d >= 3,
             "conv3d_wgrad: stride_d=%d pad_d=%d unsupported", stride_d, pad_d);
  VN_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(dy) & 15) == 0,
             "conv3d_wgrad: buffers must be 16-byte aligned");
  const int Dout = (Din + 2 * pad_d - 3) / stride_d + 1;
  CUtensorMap mx, my;
  int rc = encode_ndhwc(&mx, x, B, Din, H, W, cin, kTileH + 2);
  if (rc != VN_OK) return rc;
  rc = encode_ndhwc(&my, dy, B, Dout, H, W, 64, kTileH);
  if (rc != VN_OK) return rc;
  WgParams P{};
  P.B = B; P.Din = Din; P.Dout = Dout; P.H = H; P.W = W; P.cin = cin; P.stride_d = stride_d; P.pad_d = pad_d;
  P.tiles_h = (H + kTileH - 1) / kTileH;
  P.tiles_w = W / kTileW;
  P.n_groups = cin == 64 ? 3 : 9;
  P.n_slices = sm_count() / P.n_groups;
  if (P.n_slices < 1) P.n_slices = 1;
  P.dw = dw_taps;
  P.flags = box_flags;
  VN_CUDA(cudaFuncSetAttribute(k_conv3d_wgrad, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kWgSmem));
  k_conv3d_wgrad<<<P.n_groups * P.n_slices, kWgThreads, kWgSmem, as_stream(stream)>>>(mx, my, P);
  VN_LAUNCH_CHECK("k_conv3d_wgrad");
  return VN_OK;
}

}  // extern "C"
