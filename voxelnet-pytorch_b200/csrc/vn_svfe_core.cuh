// Per-warp evaluation of the stacked VFE for ONE voxel whose rows sit in shared memory.
// Shared by k_svfe_fused (rows staged from a materialised [K,T,7] feature buffer) and by
// k_vox_svfe_fused (rows built straight from the voxelizer's point buckets), so both paths run
// the same instruction sequence and produce identical bits.
//
// Reference semantics (voxelnet/model.py:74-82, 95-100):
//   pointwise = relu(W x + b) * s + t          Linear -> ReLU -> eval BatchNorm folded
//   agg       = max over ALL T rows (pad rows included, no mask)
//   out       = concat(pointwise, agg) * mask   mask = (max_j x_j != 0)
//   voxelwise = max over t of the masked VFE-2 output
// Rows that are bit-identical (the reference's pad rows) are passed once: max() ignores multiplicity.
#pragma once

#include <cuda_bf16.h>

#include "vn_common.cuh"

namespace vn {

struct Vfe1Params {
  float w[16 * 7];
  float b[16], s[16], t[16];
};

constexpr int kXS = 16;  // floats per VFE-1 output row kept in shared memory

// VFE-2 parameters.  A VFE-2 input row is [x1_row (16) | agg1 (16)] where agg1 is the same for every
// row of a voxel, so  W2 x + b = (b + W2[:,16:32] agg1) + W2[:,0:16] x1_row : the second half of the
// weights is applied once per voxel (read from shared memory), the first half per row (registers).
struct Vfe2Lane {  // parameters of the two output units (lane, lane + 32) owned by one lane
  float wa[16], wb[16];  // W2[unit][0..15]
  float ba, bb, sa, sb, ta, tb;
};
struct Vfe2Shared {
  float whi[16][64];     // whi[k][unit] = W2[unit][16 + k]
  float4 whi4[4][64];    // whi4[k4][unit] = W2[unit][16 + 4 k4 .. + 3]: the same values for 128-bit loads
};

__device__ __forceinline__ void load_vfe1_params(Vfe1Params& p1, const float* w1, const float* b1, const float* s1,
                                                 const float* t1) {
  for (int i = threadIdx.x; i < 16 * 7; i += blockDim.x) p1.w[i] = w1[i];
  if (threadIdx.x < 16) {
    p1.b[threadIdx.x] = b1[threadIdx.x];
    p1.s[threadIdx.x] = s1[threadIdx.x];
    p1.t[threadIdx.x] = t1[threadIdx.x];
  }
}

__device__ __forceinline__ void load_vfe2(Vfe2Lane& q, Vfe2Shared& sh, const float* w2, const float* b2, const float* s2,
                                          const float* t2, uint32_t lane) {
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    q.wa[k] = w2[lane * 32 + k];
    q.wb[k] = w2[(lane + 32) * 32 + k];
  }
  q.ba = b2[lane];  q.bb = b2[lane + 32];
  q.sa = s2[lane];  q.sb = s2[lane + 32];
  q.ta = t2[lane];  q.tb = t2[lane + 32];
  for (int i = threadIdx.x; i < 16 * 64; i += blockDim.x) {
    const float v = w2[(i & 63) * 32 + 16 + (i >> 6)];
    sh.whi[i >> 6][i & 63] = v;
    reinterpret_cast<float*>(&sh.whi4[i >> 8][i & 63])[(i >> 6) & 3] = v;
  }
}

// VFE-2 + voxel-wise max for ONE voxel.  x: the voxel's R VFE-1 output rows (kXS floats apart) in shared
// memory; agg: its 16 agg1 values.  Lane l owns units (l, l + 32).  on_bits: bit r set <=> row r has mask 1.
// A row with mask 0 is an all-zero VFE-2 input (model.py:82 masks the concat): its pointwise output is
// relu(b)*s + t, it counts for the unmasked aggregate (model.py:77) and contributes exactly 0 to the
// voxel-wise max of the masked output (model.py:100).  o[i] = voxelwise channel (lane + 32 i).
// Mask: uint64_t (up to 64 rows) or uint32_t (callers with at most 32 rows: half the bit bookkeeping per row).
template <typename Mask>
__device__ __forceinline__ int lowest_bit(Mask m);
template <>
__device__ __forceinline__ int lowest_bit<uint64_t>(uint64_t m) { return __ffsll((long long)m) - 1; }
template <>
__device__ __forceinline__ int lowest_bit<uint32_t>(uint32_t m) { return __ffs((int)m) - 1; }

template <typename Mask = uint64_t>
__device__ __forceinline__ void vfe2_voxel_max(const float* x, const float* agg, int R, Mask on_bits,
                                               const Vfe2Lane& q, const Vfe2Shared& sh, uint32_t lane, float (&o)[4]) {
  constexpr int kBits = (int)sizeof(Mask) * 8;
  const Mask evaluated = R >= kBits ? ~(Mask)0 : (Mask)(((Mask)1 << R) - (Mask)1);
  on_bits &= evaluated;
  const bool any_on = on_bits != 0;
  const bool any_off = on_bits != evaluated;
  // per-voxel part: bias + W2[:,16:32] agg1
  float ca = q.ba, cb = q.bb;
  {
    const float4* g4 = reinterpret_cast<const float4*>(agg);
#pragma unroll
    for (int k4 = 0; k4 < 4; ++k4) {
      const float4 g = g4[k4];
      const float4 wa4 = sh.whi4[k4][lane], wb4 = sh.whi4[k4][lane + 32];  // (8 x 128-bit loads per voxel instead of 32 scalar ones)
      ca = fmaf(wa4.x, g.x, ca);  cb = fmaf(wb4.x, g.x, cb);
      ca = fmaf(wa4.y, g.y, ca);  cb = fmaf(wb4.y, g.y, cb);
      ca = fmaf(wa4.z, g.z, ca);  cb = fmaf(wb4.z, g.z, cb);
      ca = fmaf(wa4.w, g.w, ca);  cb = fmaf(wb4.w, g.w, cb);
    }
  }
  float on_a = -INFINITY, on_b = -INFINITY;  // max over the rows with mask 1
  Mask todo = on_bits;
  if ((on_bits & (on_bits + 1)) == 0) {  // the rows with mask 1 are rows 0 .. n_on-1 (the usual case: real points, then the pad row)
    const int n_on = (int)(sizeof(Mask) == 8 ? __popcll((unsigned long long)on_bits) : __popc((unsigned)on_bits));
    int r = 0;
    for (; r + 1 < n_on; r += 2) {
      float a0 = ca, b0 = cb, a1 = ca, b1 = cb;
      const float4* x0 = reinterpret_cast<const float4*>(x + r * kXS);
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4) {
        const float4 v = x0[k4], w = x0[k4 + kXS / 4];
        a0 = fmaf(q.wa[4 * k4 + 0], v.x, a0);  b0 = fmaf(q.wb[4 * k4 + 0], v.x, b0);
        a1 = fmaf(q.wa[4 * k4 + 0], w.x, a1);  b1 = fmaf(q.wb[4 * k4 + 0], w.x, b1);
        a0 = fmaf(q.wa[4 * k4 + 1], v.y, a0);  b0 = fmaf(q.wb[4 * k4 + 1], v.y, b0);
        a1 = fmaf(q.wa[4 * k4 + 1], w.y, a1);  b1 = fmaf(q.wb[4 * k4 + 1], w.y, b1);
        a0 = fmaf(q.wa[4 * k4 + 2], v.z, a0);  b0 = fmaf(q.wb[4 * k4 + 2], v.z, b0);
        a1 = fmaf(q.wa[4 * k4 + 2], w.z, a1);  b1 = fmaf(q.wb[4 * k4 + 2], w.z, b1);
        a0 = fmaf(q.wa[4 * k4 + 3], v.w, a0);  b0 = fmaf(q.wb[4 * k4 + 3], v.w, b0);
        a1 = fmaf(q.wa[4 * k4 + 3], w.w, a1);  b1 = fmaf(q.wb[4 * k4 + 3], w.w, b1);
      }
      a0 = fmaf(fmaxf(a0, 0.0f), q.sa, q.ta);
      b0 = fmaf(fmaxf(b0, 0.0f), q.sb, q.tb);
      a1 = fmaf(fmaxf(a1, 0.0f), q.sa, q.ta);
      b1 = fmaf(fmaxf(b1, 0.0f), q.sb, q.tb);
      on_a = fmaxf(on_a, fmaxf(a0, a1));
      on_b = fmaxf(on_b, fmaxf(b0, b1));
    }
    todo = r < n_on ? (Mask)((Mask)1 << r) : (Mask)0;  // the odd row out takes the generic single-row path below
  }
  while (todo) {
    const int r0 = lowest_bit<Mask>(todo);
    todo &= todo - 1;
    if (todo) {  // two rows per trip: four independent FMA chains
      const int r1 = lowest_bit<Mask>(todo);
      todo &= todo - 1;
      float a0 = ca, b0 = cb, a1 = ca, b1 = cb;
      const float4* x0 = reinterpret_cast<const float4*>(x + r0 * kXS);
      const float4* x1 = reinterpret_cast<const float4*>(x + r1 * kXS);
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4) {
        const float4 v = x0[k4], w = x1[k4];
        a0 = fmaf(q.wa[4 * k4 + 0], v.x, a0);  b0 = fmaf(q.wb[4 * k4 + 0], v.x, b0);
        a1 = fmaf(q.wa[4 * k4 + 0], w.x, a1);  b1 = fmaf(q.wb[4 * k4 + 0], w.x, b1);
        a0 = fmaf(q.wa[4 * k4 + 1], v.y, a0);  b0 = fmaf(q.wb[4 * k4 + 1], v.y, b0);
        a1 = fmaf(q.wa[4 * k4 + 1], w.y, a1);  b1 = fmaf(q.wb[4 * k4 + 1], w.y, b1);
        a0 = fmaf(q.wa[4 * k4 + 2], v.z, a0);  b0 = fmaf(q.wb[4 * k4 + 2], v.z, b0);
        a1 = fmaf(q.wa[4 * k4 + 2], w.z, a1);  b1 = fmaf(q.wb[4 * k4 + 2], w.z, b1);
        a0 = fmaf(q.wa[4 * k4 + 3], v.w, a0);  b0 = fmaf(q.wb[4 * k4 + 3], v.w, b0);
        a1 = fmaf(q.wa[4 * k4 + 3], w.w, a1);  b1 = fmaf(q.wb[4 * k4 + 3], w.w, b1);
      }
      a0 = fmaf(fmaxf(a0, 0.0f), q.sa, q.ta);
      b0 = fmaf(fmaxf(b0, 0.0f), q.sb, q.tb);
      a1 = fmaf(fmaxf(a1, 0.0f), q.sa, q.ta);
      b1 = fmaf(fmaxf(b1, 0.0f), q.sb, q.tb);
      on_a = fmaxf(on_a, fmaxf(a0, a1));
      on_b = fmaxf(on_b, fmaxf(b0, b1));
    } else {
      float a = ca, b = cb;
      const float4* xr = reinterpret_cast<const float4*>(x + r0 * kXS);
#pragma unroll
      for (int k4 = 0; k4 < 4; ++k4) {
        const float4 v = xr[k4];
        a = fmaf(q.wa[4 * k4 + 0], v.x, a);  b = fmaf(q.wb[4 * k4 + 0], v.x, b);
        a = fmaf(q.wa[4 * k4 + 1], v.y, a);  b = fmaf(q.wb[4 * k4 + 1], v.y, b);
        a = fmaf(q.wa[4 * k4 + 2], v.z, a);  b = fmaf(q.wb[4 * k4 + 2], v.z, b);
        a = fmaf(q.wa[4 * k4 + 3], v.w, a);  b = fmaf(q.wb[4 * k4 + 3], v.w, b);
      }
      a = fmaf(fmaxf(a, 0.0f), q.sa, q.ta);
      b = fmaf(fmaxf(b, 0.0f), q.sb, q.tb);
      on_a = fmaxf(on_a, a);
      on_b = fmaxf(on_b, b);
    }
  }
  // unmasked aggregate: rows with mask 0 evaluate to relu(bias)*s + t
  float all_a = on_a, all_b = on_b;
  if (any_off) {
    all_a = fmaxf(all_a, fmaf(fmaxf(q.ba, 0.0f), q.sa, q.ta));
    all_b = fmaxf(all_b, fmaf(fmaxf(q.bb, 0.0f), q.sb, q.tb));
  }
  // voxelwise[c] = max_t(mask_t * out2[t][c]): rows with mask 0 contribute exactly 0
  const float floor0 = any_off ? 0.0f : -INFINITY;
  o[0] = fmaxf(any_on ? on_a : -INFINITY, floor0);
  o[1] = fmaxf(any_on ? on_b : -INFINITY, floor0);
  o[2] = fmaxf(any_on ? all_a : -INFINITY, floor0);
  o[3] = fmaxf(any_on ? all_b : -INFINITY, floor0);
}

// One voxel's 128 maxima -> voxelwise row (f32, model.py:100) and / or its cell of the resident bf16
// grid (model.py:102-106; the same round-to-nearest conversion as vn_scatter_dense_bf16).  Either
// destination may be null.
__device__ __forceinline__ void store_voxel_out(const float (&o)[4], uint32_t lane, float* vw_row, void* dense_row) {
  if (vw_row != nullptr) {
#pragma unroll
    for (int i = 0; i < 4; ++i) vw_row[lane + 32 * i] = o[i];
  }
  if (dense_row != nullptr) {  // warp-uniform
    // lane l holds channels l + 32 i; pair neighbouring lanes so that every lane stores 2 x bf16x2:
    // even lanes write groups 0 and 2, odd lanes groups 1 and 3
    const bool odd = lane & 1;
    uint32_t* dst = reinterpret_cast<uint32_t*>(dense_row);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float mine = odd ? o[2 * h + 1] : o[2 * h];       // the value of the group this lane writes
      const float give = odd ? o[2 * h] : o[2 * h + 1];       // the value the neighbour needs
      const float got = __shfl_xor_sync(0xffffffffu, give, 1);
      const __nv_bfloat162 v = odd ? __floats2bfloat162_rn(got, mine) : __floats2bfloat162_rn(mine, got);
      const int group = 2 * h + (odd ? 1 : 0);
      dst[(group * 32 + (int)(lane & ~1u)) >> 1] = *reinterpret_cast<const uint32_t*>(&v);
    }
  }
}

// rows: shared memory, 7 floats per row.  Evaluated row r (0 <= r < R) is source row
// (r < nu ? r : tlast).  x1: shared scratch of (R + 1) x kXS floats; row R receives agg1.
// o: the voxel's 128 maxima, channel (lane + 32 i) in o[i].
__device__ __forceinline__ void svfe_eval_voxel(const float* rows, int nu, int R, int tlast, float* x1,
                                                const Vfe1Params& p1, const Vfe2Lane& q, const Vfe2Shared& sh,
                                                float (&o)[4], uint32_t lane) {
  // ---- mask bits: bit r set <=> evaluated row r has max_j != 0 --------------------------------
  uint64_t on_bits = 0;
  for (int r0 = 0; r0 < R; r0 += 32) {
    const int r = r0 + lane;
    bool on = false;
    if (r < R) {
      const float* f = rows + (r < nu ? r : tlast) * 7;
      float mx = f[0];
#pragma unroll
      for (int j = 1; j < 7; ++j) mx = fmaxf(mx, f[j]);
      on = mx != 0.0f;
    }
    on_bits |= (uint64_t)__ballot_sync(0xffffffffu, on) << r0;
  }
  // ---- VFE-1: (row, unit) pairs spread over the lanes ----------------------------------------
  for (int e = lane; e < R * 16; e += 32) {
    const int r = e >> 4, u = e & 15;
    const float* f = rows + (r < nu ? r : tlast) * 7;
    float a = p1.b[u];
#pragma unroll
    for (int j = 0; j < 7; ++j) a = fmaf(p1.w[u * 7 + j], f[j], a);
    x1[r * kXS + u] = fmaf(fmaxf(a, 0.0f), p1.s[u], p1.t[u]);
  }
  __syncwarp();
  if (lane < 16) {  // agg1[u] = max over evaluated rows (== max over all T rows, masked or not)
    float a = x1[lane];
    for (int r = 1; r < R; ++r) a = fmaxf(a, x1[r * kXS + lane]);
    x1[R * kXS + lane] = a;
  }
  __syncwarp();
  vfe2_voxel_max(x1, x1 + R * kXS, R, on_bits, q, sh, lane, o);
}

}  // namespace vn
