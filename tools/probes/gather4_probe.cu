// Probe (GPU box): semantics of cp.async.bulk.tensor.2d ... tile::gather4 on sm_100a.
//   * which box shape must the 2-D tensor map carry ({64,1} or {64,4})?
//   * are out-of-range row indices zero-filled, and do they count toward the mbarrier's tx bytes?
//   * is SWIZZLE_128B applied by absolute shared-memory address (so that two gathers fill one 8-row atom)?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I include -I voxelnet-pytorch_b200/csrc tools/probes/gather4_probe.cu -o tools/probes/gather4_probe
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda.h>
#include <cuda_runtime.h>
#include "vn_common.cuh"
using namespace vn;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ void tma_gather4(void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int r0, int r1, int r2, int r3) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      :
      : "r"(smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(bar)), "r"(c0), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
      : "memory");
}

// out: [0] = status (1 ok, 0 timeout), [1] = clocks; smem_out: 2048 bytes (two atoms)
__global__ void k_probe(const __grid_constant__ CUtensorMap tm, int4 rows_a, int4 rows_b, int c0, uint32_t tx, long long* out,
                        uint8_t* smem_out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  for (int i = threadIdx.x; i < 2048 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0xEEEEEEEEu;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); fence_proxy_async_smem(); }
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    mbar_arrive_expect_tx(&bar, tx);
    tma_gather4(smem, &tm, &bar, c0, rows_a.x, rows_a.y, rows_a.z, rows_a.w);
    tma_gather4(smem + 512, &tm, &bar, c0, rows_b.x, rows_b.y, rows_b.z, rows_b.w);
    int ok = 0;
    while (clock64() - t0 < 20000000LL) {
      if (mbar_test_wait(&bar, 0)) { ok = 1; break; }
    }
    out[0] = ok;
    out[1] = clock64() - t0;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) smem_out[i] = smem[i];
}

__device__ __forceinline__ bool bounded_wait(uint64_t* bar, uint32_t parity) {
  const long long t0 = clock64();
  while (!mbar_test_wait(bar, parity))
    if (clock64() - t0 > 50000000LL) return false;
  return true;
}

// throughput: every CTA issues `n` gather4 (4 x 128 B) per round from 32 lanes, `rounds` rounds, rows pseudo-random
// (frac_oob of them out of range)
__global__ void __launch_bounds__(64, 1) k_rate(const __grid_constant__ CUtensorMap tm, int R, int rounds, int oob_every, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar[4];
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1); mbar_fence_init(); fence_proxy_async_smem(); }
  __syncthreads();
  if (threadIdx.x < 32) {
    const uint32_t lane = threadIdx.x;
    uint32_t seed = blockIdx.x * 977u + lane * 31u + 1u;
    const long long t0 = clock64();
    for (int r = 0; r < rounds; ++r) {
      const int s = r & 3;
      if (r >= 4 && !bounded_wait(&bar[s], ((r / 4) - 1) & 1)) { if (lane == 0) out[blockIdx.x] = -1; return; }
      if (lane == 0) mbar_arrive_expect_tx(&bar[s], 36 * 512);
      __syncwarp();
      for (int q = lane; q < 36; q += 32) {
        int rr[4];
        for (int j = 0; j < 4; ++j) {
          seed = seed * 1664525u + 1013904223u;
          const bool oob = oob_every > 0 && ((seed >> 8) % (uint32_t)oob_every) != 0;
          rr[j] = oob ? R + 7 : (int)((seed >> 10) % (uint32_t)R);
        }
        tma_gather4(smem + s * 18432 + q * 512, &tm, &bar[s], 0, rr[0], rr[1], rr[2], rr[3]);
      }
    }
    for (int r = rounds; r < rounds + 4; ++r)
      if (r >= 4 && !bounded_wait(&bar[r & 3], ((r / 4) - 1) & 1)) { if (lane == 0) out[blockIdx.x] = -1; return; }
    if (lane == 0) out[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
    printf("no cuTensorMapEncodeTiled\n");
    return 1;
  }
  EncodeTiledFn encode = reinterpret_cast<EncodeTiledFn>(p);
  const int R = 64;
  uint16_t* h = (uint16_t*)malloc(R * 128 * 2);
  for (int r = 0; r < R; ++r)
    for (int c = 0; c < 128; ++c) h[r * 128 + c] = (uint16_t)(r * 256 + c);
  uint16_t* d;
  cudaMalloc(&d, R * 128 * 2);
  cudaMemcpy(d, h, R * 128 * 2, cudaMemcpyHostToDevice);
  long long* dout;
  uint8_t* dsm;
  cudaMalloc(&dout, 148 * 8 * 2);
  cudaMalloc(&dsm, 2048);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 8192);
  for (int variant = 0; variant < 2; ++variant) {
    CUtensorMap tm;
    const cuuint64_t gdim[2] = {128, (cuuint64_t)R};
    const cuuint64_t gstr[1] = {256};
    const cuuint32_t box[2] = {64, variant == 0 ? 1u : 4u};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult cr = encode(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("== variant box {64,%d}: encode -> %d\n", variant == 0 ? 1 : 4, (int)cr);
    if (cr != CUDA_SUCCESS) continue;
    // case A: all rows in range; case B: one out of range (+), one negative
    for (int cs = 0; cs < 3; ++cs) {
      const int4 ra = cs == 0 ? make_int4(5, 9, 17, 2) : (cs == 1 ? make_int4(5, R + 3, 17, 2) : make_int4(5, -1, 17, 2));
      const int4 rb = make_int4(40, 41, 63, 0);
      cudaMemset(dout, 0, 16);
      k_probe<<<1, 128, 8192>>>(tm, ra, rb, 64, 1024u, dout, dsm);
      cudaError_t e = cudaDeviceSynchronize();
      long long ho[2];
      uint8_t hs[2048];
      cudaMemcpy(ho, dout, 16, cudaMemcpyDeviceToHost);
      cudaMemcpy(hs, dsm, 2048, cudaMemcpyDeviceToHost);
      printf("case %d rows_a={%d,%d,%d,%d}: err=%s completed=%lld clk=%lld\n", cs, ra.x, ra.y, ra.z, ra.w, cudaGetErrorString(e), ho[0], ho[1]);
      if (e != cudaSuccess) return 2;
      // decode: smem row j (0..7) of the atom, 16-byte chunk k: expect table row rows[j], columns 64 + 8*(k ^ (j & 7)) ..
      const int rows8[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
      int bad = 0;
      for (int j = 0; j < 8; ++j) {
        const uint16_t* row = reinterpret_cast<const uint16_t*>(hs + j * 128);
        printf("  smem row %d (table row %d): ", j, rows8[j]);
        for (int k = 0; k < 8; ++k) printf("%04x ", row[k * 8]);
        for (int k = 0; k < 8; ++k) {
          const int src_chunk = k ^ (j & 7);
          const bool inr = rows8[j] >= 0 && rows8[j] < R;
          const uint16_t want = inr ? (uint16_t)(rows8[j] * 256 + 64 + src_chunk * 8) : 0;
          if (row[k * 8] != want) ++bad;
        }
        printf("\n");
      }
      printf("  -> %s (second KB first word %08x)\n", bad == 0 ? "MATCHES swizzle-by-address, OOB rows zero" : "MISMATCH", *reinterpret_cast<uint32_t*>(hs + 1024));
    }
    // rate
    cudaFuncSetAttribute(k_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 18432 + 1024);
    for (int oob_every = 0; oob_every <= 12; oob_every += 12) {
      const int rounds = 2000;
      k_rate<<<148, 64, 4 * 18432 + 1024>>>(tm, R, rounds, oob_every, dout);
      cudaError_t e = cudaDeviceSynchronize();
      long long ho[148];
      cudaMemcpy(ho, dout, 148 * 8, cudaMemcpyDeviceToHost);
      long long mx = 0;
      for (int i = 0; i < 148; ++i) { if (ho[i] < 0) { printf("CTA %d timed out\n", i); break; } mx = ho[i] > mx ? ho[i] : mx; }
      printf("rate (%s): err=%s, %d boxes of 36 gather4 per CTA, 148 CTAs: %.1f clk per box (max CTA)\n",
             oob_every ? "11/12 rows out of range" : "all rows in range (64-row table)", cudaGetErrorString(e), rounds, (double)mx / rounds);
    }
  }
  return 0;
}
