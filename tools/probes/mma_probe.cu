// Probe (GPU box): how long does ONE tcgen05.mma.cta_group::1.kind::f16 of shape 128 x N x 16 take when issued back to
// back from shared-memory operands (SWIZZLE_128B, K-major)?  N = 64 / 128 / 192 / 256, every SM busy.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I include -I voxelnet-pytorch_b200/csrc tools/probes/mma_probe.cu -o tools/probes/mma_probe
#include <cstdio>
#include <cuda_runtime.h>
#include "vn_common.cuh"
using namespace vn;

__global__ void __launch_bounds__(128, 1) k_probe(int N, int rounds, int a_stride16, long long* out, int mode) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint64_t bars[4];
  __shared__ uint4 msgs[4];
  __shared__ uint32_t tmem_base_s;
  const uint32_t warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (64 + 32) * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); for (int i = 0; i < 4; ++i) { mbar_init(&bars[i], 1); msgs[i] = make_uint4(i, 0, 0, 0); } mbar_fence_init(); fence_proxy_async_smem(); }
  if (warp == 0) { tmem_alloc(&tmem_base_s, 512); tmem_relinquish(); }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem = tmem_base_s;
  if (warp == 1) {
    const uint32_t a_lo0 = ((smem_u32(smem) >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t b_lo0 = ((smem_u32(smem + 64 * 1024) >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t hi = (uint64_t)0x40004040u << 32;
    long long t0 = 0, t1 = 0;
    if (mode == 0) {
      if (elect_one()) {
        t0 = clock64();
        for (int r = 0; r < rounds; ++r) {
          const uint32_t a = a_lo0 + (r & 3) * a_stride16;  // rotate over 4 A tiles (like the conv's ring)
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem, hi | (a + k * 2), hi | (b_lo0 + k * 2), idesc, 1u);
        }
        umma_commit(&bar);
      }
    } else if (mode == 1) {  // a commit (to a barrier nobody waits on) after every 12 MMAs
      if (elect_one()) {
        t0 = clock64();
        for (int r = 0; r < rounds; ++r) {
          const uint32_t a = a_lo0 + (r & 3) * a_stride16;
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem, hi | (a + k * 2), hi | (b_lo0 + k * 2), idesc, 1u);
          if (r % 3 == 2) umma_commit(&bars[(r / 3) & 3]);
        }
        umma_commit(&bar);
      }
    } else if (mode >= 100) {
      // switch the destination accumulator every `period` MMAs (mode = 100 + period), rotating over 4 accumulators
      const int sh = mode - 100;  // period = 1 << sh
      if (elect_one()) {
        t0 = clock64();
        int cnt = 0;
        for (int r = 0; r < rounds; ++r) {
          const uint32_t a = a_lo0 + (r & 3) * a_stride16;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const uint32_t d = tmem + ((cnt >> sh) & 3) * 128;
            umma_bf16_ss(d, hi | (a + k * 2), hi | (b_lo0 + k * 2), idesc, 1u);
            ++cnt;
          }
        }
        umma_commit(&bar);
      }
    } else if (mode >= 4) {
      // single-thread issuer as in k_conv3d_bn_relu: per "message" 12 MMAs + commit; mode bits: 1 = tcgen05.fence::after_thread_sync
      // per message, 2 = barrier probe + message read between the MMA groups, 4 (mode 8+) = first MMA overwrites + rotating accumulators
      const int opt = mode - 4;
      if (elect_one()) {
        t0 = clock64();
        uint32_t acc = 0;
        for (int r = 0; r < rounds; r += 3) {
          const int m = r / 3;
          const uint32_t a = a_lo0 + (m & 3) * a_stride16;
          const uint32_t d = (opt & 4) ? tmem + ((m * 3) & 7) * 64 % (512 - N) : tmem;
          if (opt & 1) tc_fence_after_sync();
          umma_bf16_ss(d, hi | a, hi | b_lo0, idesc, (opt & 8) ? 0u : 1u);
#pragma unroll
          for (int k = 1; k < 4; ++k) umma_bf16_ss(d, hi | (a + k * 2), hi | (b_lo0 + k * 2), idesc, 1u);
          if (opt & 2) {
            const bool ok = mbar_test_wait(&bars[m & 3], 1);
            const uint4 mm = lds_u4(&msgs[m & 3]);
            acc += mm.x + (ok ? 1 : 0);
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) umma_bf16_ss(d, hi | (a + 64 + k * 2), hi | (b_lo0 + 512 + k * 2), idesc, 1u);
          umma_commit(&bars[m & 3]);
        }
        umma_commit(&bar);
        if (acc == 0xdeadbeef) out[200] = acc;
      }
    } else {  // the conv issuer's structure: per 12 MMAs a warp-wide barrier probe + message read + three elect sections
      t0 = clock64();
      uint32_t acc = 0;
      for (int r = 0; r < rounds; r += 3) {
        const uint32_t a = a_lo0 + ((r / 3) & 3) * a_stride16;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) umma_bf16_ss(tmem, hi | (a + k * 2), hi | (b_lo0 + k * 2), idesc, 1u);
        }
        __syncwarp();
        const bool ok = __all_sync(0xffffffffu, mbar_test_wait(&bars[(r / 3) & 3], 1));
        const uint4 m = lds_u4(&msgs[(r / 3) & 3]);
        acc += m.x + (ok ? 1 : 0);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 8; ++k) umma_bf16_ss(tmem, hi | (a + 64 + k * 2), hi | (b_lo0 + 512 + k * 2), idesc, 1u);
          if (mode == 2) umma_commit(&bars[(r / 3) & 3]);
        }
        __syncwarp();
      }
      if (elect_one()) umma_commit(&bar);
      if (acc == 0xdeadbeef) out[200] = acc;
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    t1 = clock64();
    if (lane_id() == 0) out[blockIdx.x] = t1 - t0;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tmem, 512); }
}

int main() {
  long long* out;
  cudaMalloc(&out, 256 * 8);
  cudaFuncSetAttribute(k_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int rounds = 2000;
  for (int mode : {100, 101, 102, 103, 104, 106, 120})
  for (int grid : {148}) {
    for (int N : {64, 128}) {  // (N = 128 at rotating offsets 0,128,256,384)
      k_probe<<<grid, 128, 98 * 1024 + 1024>>>(N, rounds - rounds % 3, 1152, out, mode);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      long long h[148];
      cudaMemcpy(h, out, grid * 8, cudaMemcpyDeviceToHost);
      double s = 0;
      for (int i = 0; i < grid; ++i) s += h[i];
      if (mode == 4 && N == 64) printf("mode 4 = same D, accumulate; 8 = rotating D, accumulate; 12 = same D, first MMA of 12 overwrites; 16 = rotating D + overwrite\n");
      printf("mode %d grid %3d  N=%3d  clk per MMA = %.1f  (pipe floor %d)\n", mode, grid, N, s / grid / ((rounds - rounds % 3) * 4.0), 128 * N / 256);
    }
  }
  return 0;
}
