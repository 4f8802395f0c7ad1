// Batched GPU voxelizer: the replacement for pcl_to_voxels (reference voxelnet/utils.py:10-100)
// plus the batch column of prepare_voxel (voxelnet/dataset.py:101-119).
//
// Design (see DESIGN.md "Voxelizer"): the grid is small and dense (1.4 M cells for Car), so the
// sorted-unique step (np.unique(axis=0), utils.py:63) is done WITHOUT a sort: an occupancy bitmap per
// frame (176 KB, L2 resident) plus a prefix sum of its popcounts gives every occupied cell its
// rank in ascending (z, y, x) order - exactly np.unique's lexicographic order.  The first-T-points
// rule (utils.py:78-84) needs a deterministic order that plain atomics cannot give, so points
// are first bucketed per voxel in arrival order (warp-aggregated atomics) and each voxel's short
// list is then ordered by original point index when its feature rows are written.
//
//   P0 zero            bitmap + per-voxel counters
//   P1 mark            float4 point load, IEEE add/div/floor, range crop, cell key, bitmap bit
//   P2 scan words      per frame: exclusive prefix of popcounts  -> K per frame
//   P3 frame offsets   exclusive prefix of K over frames         -> voxel_offsets
//   P4 count           key -> voxel row; arrival slot via warp-aggregated atomicAdd
//   P5 segment scan    per frame: exclusive prefix of counts     -> CSR segment starts
//   P6 place           point index -> its voxel's segment
//   P7 features        per voxel (one warp): T smallest indices ascending, gather, fp32 sequential
//                      sum, fp64 centroid, write [T,7] row block, coords, count
#include "vn_svfe_core.cuh"

namespace vn {
namespace {

constexpr uint32_t kInvalidKey = 0xFFFFFFFFu;
constexpr int kCtrItems = 0, kCtrBigCount = 1, kCtrBigNext = 2;  // VoxWorkspace::counters slots
constexpr int kMaxT = 64;  // smem staging rows per warp (T is 35 or 45)
constexpr int kScanChunk = 1024;  // elements per block of the two-level scans (256 threads x 4)

struct VoxWorkspace {
  uint32_t* bitmap;      // [n_frames * words]   occupancy bits, cell = (z*H + y)*W + x
  uint32_t* wordprefix;  // [n_frames * words]   exclusive popcount prefix within the frame
  int32_t* frame_k;      // [n_frames]           voxels per frame
  int32_t* cnt;          // [n_points]           points per voxel row (uncapped)
  int32_t* counters;     // [64]                 work counters of the persistent kernels (zeroed with the bitmap)
  int32_t* keyrow;       // [n_points]           P1: cell key; P4 onwards: voxel row (-1 = cropped)
  int32_t* arrival;      // [n_points]           arrival slot inside the voxel
  int32_t* seg;          // [n_points]           segment start of each voxel row in pidx
  int32_t* pidx;         // [n_points]           point indices bucketed per voxel
  int32_t* biglist;      // [n_points / (min(31,T)+1) + 1]  voxel rows with more than 31 (or more than T) points
  uint32_t* wordpart;    // [n_frames * word_chunks]  popcount of each 1024-word chunk of the bitmap (two-level scan)
  int32_t* segpart;      // [n_frames * seg_chunks]   points in each 1024-row chunk of a frame's voxel rows
  size_t word_chunks, seg_chunks;
  size_t words;          // bitmap words per frame
  size_t zero_bytes;     // bitmap..cnt are contiguous: one fill
  size_t total_bytes;
};

__host__ VoxWorkspace carve(const vn_grid_t& g, int n_frames, int64_t n_points, void* base) {
  VoxWorkspace w{};
  const size_t cells = (size_t)g.D * g.H * g.W;
  w.words = align_up((cells + 31) / 32, 32);
  const size_t np = align_up((size_t)(n_points > 0 ? n_points : 1), 64);
  char* p = static_cast<char*>(base);
  size_t off = 0;
  auto take = [&](size_t bytes) {
    char* r = p ? p + off : nullptr;
    off += align_up(bytes, 256);
    return r;
  };
  w.bitmap = reinterpret_cast<uint32_t*>(take((size_t)n_frames * w.words * 4));
  w.cnt = reinterpret_cast<int32_t*>(take(np * 4));
  w.counters = reinterpret_cast<int32_t*>(take(64 * 4));
  w.zero_bytes = off;
  w.wordprefix = reinterpret_cast<uint32_t*>(take((size_t)n_frames * w.words * 4));
  w.frame_k = reinterpret_cast<int32_t*>(take((size_t)n_frames * 4));
  w.keyrow = reinterpret_cast<int32_t*>(take(np * 4));
  w.arrival = reinterpret_cast<int32_t*>(take(np * 4));
  w.seg = reinterpret_cast<int32_t*>(take(np * 4));
  w.pidx = reinterpret_cast<int32_t*>(take(np * 4));
  w.biglist = reinterpret_cast<int32_t*>(take((np / (size_t)((g.T < 31 ? g.T : 31) + 1) + 1) * 4));
  w.word_chunks = (w.words + kScanChunk - 1) / kScanChunk;
  w.seg_chunks = ((cells < np ? cells : np) + kScanChunk - 1) / kScanChunk;  // a frame has at most min(cells, points) voxels
  w.wordpart = reinterpret_cast<uint32_t*>(take((size_t)n_frames * w.word_chunks * 4));
  w.segpart = reinterpret_cast<int32_t*>(take((size_t)n_frames * w.seg_chunks * 4));
  w.total_bytes = off;
  return w;
}

// utils.py:37-57.  shifted = p + lidar_coord (fp32), idx = floor(shifted / voxel_size) with a true
// IEEE fp32 division (a reciprocal multiply moves ~1e-3 of the points to a neighbouring cell,
// SURVEY F2).  The bounds test is done on the floored float, which equals the reference's test on
// the int64 cast for every finite value and rejects NaN/inf like the cast's INT64_MIN does.
__device__ __forceinline__ bool point_cell(const float4& p, const vn_grid_t& g, int& z, int& y, int& x) {
  const float fx = floorf(__fdiv_rn(__fadd_rn(p.x, g.off_x), g.vs_x));
  const float fy = floorf(__fdiv_rn(__fadd_rn(p.y, g.off_y), g.vs_y));
  const float fz = floorf(__fdiv_rn(__fadd_rn(p.z, g.off_z), g.vs_z));
  const bool ok = (fx >= 0.0f) & (fx < (float)g.W) & (fy >= 0.0f) & (fy < (float)g.H) & (fz >= 0.0f) &
                  (fz < (float)g.D);
  x = (int)fx;
  y = (int)fy;
  z = (int)fz;
  return ok;
}

__global__ void k_fill_zero16(uint4* __restrict__ p, size_t n16) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride)
    p[i] = make_uint4(0, 0, 0, 0);
}

// P1: one thread per point; blockIdx.y = frame.
__global__ void __launch_bounds__(256) k_vox_mark(const vn_grid_t g, const float4* __restrict__ points,
                                                   const int32_t* __restrict__ frame_offsets,
                                                   uint32_t* __restrict__ bitmap, size_t words,
                                                   int32_t* __restrict__ keyrow) {
  const int f = blockIdx.y;
  const int beg = frame_offsets[f], end = frame_offsets[f + 1];
  uint32_t* bm = bitmap + (size_t)f * words;
  for (int i = beg + blockIdx.x * blockDim.x + threadIdx.x; i < end; i += gridDim.x * blockDim.x) {
    const float4 p = ld_stream_f4(points + i);
    int z, y, x;
    uint32_t key = kInvalidKey;
    if (point_cell(p, g, z, y, x)) {
      key = (uint32_t)((z * g.H + y) * g.W + x);
      const uint32_t bit = 1u << (key & 31);
      uint32_t* wptr = bm + (key >> 5);
      // most points land in a voxel whose bit is already set: test before paying for the atomic
      if ((*reinterpret_cast<volatile uint32_t*>(wptr) & bit) == 0) atomicOr(wptr, bit);
    }
    keyrow[i] = (int32_t)key;
  }
}

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, uint32_t lane) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= (uint32_t)d) v += n;
  }
  return v;
}

// Block-wide exclusive scan of one value per thread (256 threads); returns the prefix, *total = block sum.
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t* total) {
  __shared__ uint32_t warp_sum[8];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t incl = warp_incl_scan(v, lane);
  __syncthreads();  // (warp_sum may still be read from the previous call)
  if (lane == 31) warp_sum[warp] = incl;
  __syncthreads();
  uint32_t base = 0, all = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint32_t t = warp_sum[i];
    if (i < (int)warp) base += t;
    all += t;
  }
  *total = all;
  return base + incl - v;
}

// P2, two levels (a frame's 44,032 bitmap words used to be scanned by ONE block: 16 blocks on 148 SMs).
// a: popcount of every 1024-word chunk.  b: chunk base = sum of the chunks before it, then the scan inside the chunk;
// the block of a frame's last chunk also writes the frame's voxel count.
__global__ void __launch_bounds__(256) k_vox_word_partials(const uint32_t* __restrict__ bitmap, size_t words, int chunks,
                                                            uint32_t* __restrict__ wordpart) {
  const int f = blockIdx.y, c = blockIdx.x;
  const uint32_t* bm = bitmap + (size_t)f * words + (size_t)c * kScanChunk;
  const size_t left = words - (size_t)c * kScanChunk;
  uint32_t sum = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const size_t i = (size_t)threadIdx.x * 4 + j;
    if (i < left) sum += __popc(bm[i]);
  }
  uint32_t total;
  block_excl_scan_256(sum, &total);
  if (threadIdx.x == 0) wordpart[(size_t)f * chunks + c] = total;
}

__global__ void __launch_bounds__(256) k_vox_word_scan(const uint32_t* __restrict__ bitmap, size_t words, int chunks,
                                                        const uint32_t* __restrict__ wordpart,
                                                        uint32_t* __restrict__ wordprefix, int32_t* __restrict__ frame_k) {
  __shared__ uint32_t s_base;
  const int f = blockIdx.y, c = blockIdx.x;
  if (threadIdx.x < 32) {
    uint32_t b = 0;
    for (int i = threadIdx.x; i < c; i += 32) b += wordpart[(size_t)f * chunks + i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) b += __shfl_xor_sync(0xffffffffu, b, d);
    if (threadIdx.x == 0) s_base = b;
  }
  __syncthreads();
  const uint32_t* bm = bitmap + (size_t)f * words + (size_t)c * kScanChunk;
  uint32_t* wp = wordprefix + (size_t)f * words + (size_t)c * kScanChunk;
  const size_t left = words - (size_t)c * kScanChunk;
  uint32_t pc[4], sum = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const size_t i = (size_t)threadIdx.x * 4 + j;
    pc[j] = i < left ? __popc(bm[i]) : 0;
    sum += pc[j];
  }
  uint32_t total;
  uint32_t run = s_base + block_excl_scan_256(sum, &total);
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const size_t i = (size_t)threadIdx.x * 4 + j;
    if (i < left) wp[i] = run;
    run += pc[j];
  }
  if (c == chunks - 1 && threadIdx.x == 0) frame_k[f] = (int32_t)(s_base + total);
}

// P3: exclusive prefix over frames (one warp).
__global__ void k_vox_frame_offsets(const int32_t* __restrict__ frame_k, int n_frames,
                                    int32_t* __restrict__ voxel_offsets) {
  const uint32_t lane = threadIdx.x;
  uint32_t run = 0;
  for (int f0 = 0; f0 < n_frames; f0 += 32) {
    const int f = f0 + lane;
    const uint32_t k = f < n_frames ? (uint32_t)frame_k[f] : 0;
    const uint32_t incl = warp_incl_scan(k, lane);
    if (f < n_frames) voxel_offsets[f] = (int32_t)(run + incl - k);
    run += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (lane == 0) voxel_offsets[n_frames] = (int32_t)run;
}

// P4: key -> voxel row, arrival slot.  Lanes of a warp that hit the same voxel share one atomic.
__global__ void __launch_bounds__(256) k_vox_count(const int32_t* __restrict__ frame_offsets,
                                                    const uint32_t* __restrict__ bitmap,
                                                    const uint32_t* __restrict__ wordprefix, size_t words,
                                                    const int32_t* __restrict__ voxel_offsets,
                                                    int32_t* __restrict__ keyrow, int32_t* __restrict__ arrival,
                                                    int32_t* __restrict__ cnt) {
  const int f = blockIdx.y;
  const int beg = frame_offsets[f], end = frame_offsets[f + 1];
  const uint32_t* bm = bitmap + (size_t)f * words;
  const uint32_t* wp = wordprefix + (size_t)f * words;
  const int row0 = voxel_offsets[f];
  const uint32_t lane = threadIdx.x & 31;
  const int span = gridDim.x * blockDim.x;
  // warp-uniform trip count so that __match_any_sync sees whole warps
  for (int base = beg + blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < end; base += span) {
    const int i = base + lane;
    int row = -1;
    if (i < end) {
      const uint32_t key = (uint32_t)keyrow[i];
      if (key != kInvalidKey) {
        const uint32_t w = key >> 5, b = key & 31;
        row = row0 + (int)(wp[w] + __popc(bm[w] & ((1u << b) - 1u)));
      }
    }
    const uint32_t peers = __match_any_sync(0xffffffffu, row);
    if (row >= 0) {
      const int leader = __ffs(peers) - 1;
      int slot0 = 0;
      if ((int)lane == leader) slot0 = atomicAdd(cnt + row, __popc(peers));
      slot0 = __shfl_sync(peers, slot0, leader);
      arrival[i] = slot0 + __popc(peers & ((1u << lane) - 1u));
    }
    if (i < end) keyrow[i] = row;
  }
}

// P5, two levels: seg[row] = frame's first point + exclusive prefix of cnt over the frame's voxel rows.  Up to 64 blocks per
// frame, each taking the frame's 1024-row chunks c = blockIdx.x, + gridDim.x, ... (a KITTI frame has ~12 chunks).
// a: points per chunk.  b: chunk base from the partials, scan inside the chunk, and the list of the voxels the packed SVFE
// kernel handles one per warp (heavy items, scheduled first).
__global__ void __launch_bounds__(256) k_vox_seg_partials(const int32_t* __restrict__ voxel_offsets, const int32_t* __restrict__ cnt,
                                                           int chunks, int32_t* __restrict__ segpart) {
  const int f = blockIdx.y;
  const int r0 = voxel_offsets[f], r1 = voxel_offsets[f + 1];
  const int nch = (r1 - r0 + kScanChunk - 1) / kScanChunk;
  for (int c = blockIdx.x; c < nch && c < chunks; c += gridDim.x) {
    uint32_t sum = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = r0 + c * kScanChunk + (int)threadIdx.x * 4 + j;
      if (r < r1) sum += (uint32_t)cnt[r];
    }
    uint32_t total;
    block_excl_scan_256(sum, &total);
    if (threadIdx.x == 0) segpart[(size_t)f * chunks + c] = (int32_t)total;
  }
}

__global__ void __launch_bounds__(256) k_vox_segment_scan(const int32_t* __restrict__ frame_offsets,
                                                           const int32_t* __restrict__ voxel_offsets,
                                                           const int32_t* __restrict__ cnt, int chunks,
                                                           const int32_t* __restrict__ segpart,
                                                           int32_t* __restrict__ seg, int32_t* __restrict__ counters,
                                                           int32_t* __restrict__ biglist, int T) {
  __shared__ uint32_t s_base;
  const int f = blockIdx.y;
  const int r0 = voxel_offsets[f], r1 = voxel_offsets[f + 1];
  const int nch = (r1 - r0 + kScanChunk - 1) / kScanChunk;
  for (int c = blockIdx.x; c < nch && c < chunks; c += gridDim.x) {
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t b = 0;
      for (int i = threadIdx.x; i < c; i += 32) b += (uint32_t)segpart[(size_t)f * chunks + i];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) b += __shfl_xor_sync(0xffffffffu, b, d);
      if (threadIdx.x == 0) s_base = (uint32_t)frame_offsets[f] + b;
    }
    __syncthreads();
    uint32_t v[4], sum = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = r0 + c * kScanChunk + (int)threadIdx.x * 4 + j;
      v[j] = r < r1 ? (uint32_t)cnt[r] : 0u;
      if (v[j] > 31u || v[j] > (uint32_t)T) biglist[atomicAdd(&counters[kCtrBigCount], 1)] = r;
      sum += v[j];
    }
    uint32_t total;
    uint32_t run = s_base + block_excl_scan_256(sum, &total);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = r0 + c * kScanChunk + (int)threadIdx.x * 4 + j;
      if (r < r1) seg[r] = (int32_t)run;
      run += v[j];
    }
  }
}

// P6
__global__ void __launch_bounds__(256) k_vox_place(int64_t n_points, const int32_t* __restrict__ keyrow,
                                                    const int32_t* __restrict__ arrival,
                                                    const int32_t* __restrict__ seg, int32_t* __restrict__ pidx) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_points; i += stride) {
    const int row = keyrow[i];
    if (row >= 0) pidx[seg[row] + arrival[i]] = (int32_t)i;
  }
}

__device__ __forceinline__ int warp_min_i32(int v) {
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, d));
  return v;
}

__device__ __forceinline__ float f4_comp(const float4& v, int j) {
  return j == 0 ? v.x : (j == 1 ? v.y : (j == 2 ? v.z : v.w));
}

// ---- per-voxel point selection, shared by P7 and the fused voxel->SVFE kernel -----------------
// Loads the voxel's bucket, orders it by original point index and leaves the first m = min(n, T)
// points in pts[0..m) (shared memory, one float4 per slot).  Returns m.
// scratch (optional): shared memory for scratch_cap ints; lets a voxel with 32 < n <= scratch_cap points be
// ordered by rank counting instead of m passes over its bucket.
__device__ __forceinline__ int warp_select_points(const float4* __restrict__ points, const int32_t* __restrict__ pidx,
                                                  int n, int s0, int T, float4* pts, uint32_t lane,
                                                  int* scratch = nullptr, int scratch_cap = 0) {
  const int m = n < T ? n : T;
  if (n <= 32) {
    // rank of every index among the voxel's indices == its slot
    const int idx = (int)lane < n ? pidx[s0 + lane] : 0x7fffffff;
    int rank = 0;
    for (int j = 0; j < n; ++j) rank += (__shfl_sync(0xffffffffu, idx, j) < idx) ? 1 : 0;
    if ((int)lane < n && rank < m) pts[rank] = points[idx];
  } else if (n <= scratch_cap) {
    for (int j = lane; j < n; j += 32) scratch[j] = pidx[s0 + j];
    __syncwarp();
    for (int e = lane; e < n; e += 32) {
      const int idx = scratch[e];
      int rank = 0;
      for (int j = 0; j < n; ++j) rank += scratch[j] < idx ? 1 : 0;
      if (rank < m) pts[rank] = points[idx];
    }
  } else {
    // rare: more than a warp's worth of points.  Select the m smallest indices in ascending
    // order by repeated min-above-threshold over the segment.
    int last = -1, mine0 = -1, mine1 = -1;
    for (int s = 0; s < m; ++s) {
      int cand = 0x7fffffff;
      for (int j = lane; j < n; j += 32) {
        const int v = pidx[s0 + j];
        if (v > last && v < cand) cand = v;
      }
      cand = warp_min_i32(cand);
      last = cand;
      if ((int)lane == (s & 31)) {
        if (s < 32) mine0 = cand; else mine1 = cand;
      }
    }
    if (mine0 >= 0) pts[lane] = points[mine0];
    if (mine1 >= 0) pts[32 + lane] = points[mine1];
  }
  __syncwarp();
  return m;
}

// utils.py:87-88: fp32 sum over ALL T slots in slot order (numpy reduces axis 1 sequentially), the
// zero pads included - they only matter for the sign of a -0.0 sum; float32 / int64 promotes to
// float64.  Every lane returns the three means.
__device__ __forceinline__ void warp_centroid(const float4* pts, int m, int T, uint32_t lane, double& mx, double& my,
                                              double& mz) {
  double mean = 0.0;
  if (lane < 3) {
    float s = f4_comp(pts[0], lane);
    for (int t = 1; t < m; ++t) s = __fadd_rn(s, f4_comp(pts[t], lane));
    if (m < T) s = __fadd_rn(s, 0.0f);
    mean = (double)s / (double)m;
  }
  mx = __shfl_sync(0xffffffffu, mean, 0);
  my = __shfl_sync(0xffffffffu, mean, 1);
  mz = __shfl_sync(0xffffffffu, mean, 2);
}

// feature_buffer[row, t, j] for t < m (real point) or a pad slot (utils.py:71, 87-88).
__device__ __forceinline__ float feature_value(const float4* pts, int m, int t, int j, double mx, double my, double mz) {
  if (j < 4) return t < m ? f4_comp(pts[t], j) : 0.0f;
  const float b = t < m ? f4_comp(pts[t], j - 4) : 0.0f;
  const double c = j == 4 ? mx : (j == 5 ? my : mz);
  return __double2float_rn((double)b - c);  // float64 subtraction, rounded on store
}

// Returns the voxel's cell index ((b*D + z)*H + y)*W + x in the batched dense grid.
__device__ __forceinline__ int64_t write_coords(const vn_grid_t& g, const float4& p0, int n_frames,
                                                const int32_t* __restrict__ voxel_offsets, int64_t row,
                                                long long* __restrict__ coords) {
  int z, y, x;
  point_cell(p0, g, z, y, x);
  int lo = 0, hi = n_frames;  // frame of this row: last f with voxel_offsets[f] <= row
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if ((int64_t)voxel_offsets[mid] <= row) lo = mid; else hi = mid;
  }
  if (coords != nullptr) {  // null: k_vox_coords already wrote them and a side stream may be reading
    longlong2* c = reinterpret_cast<longlong2*>(coords + row * 4);
    c[0] = make_longlong2(lo, z);
    c[1] = make_longlong2(y, x);
  }
  return (((int64_t)lo * g.D + z) * g.H + y) * g.W + x;
}

// P7: one warp per voxel row -> materialised [T,7] feature rows (the reference's voxel dict).
constexpr int kFeatWarps = 8;
__global__ void __launch_bounds__(kFeatWarps * 32) k_vox_features(
    const vn_grid_t g, const float4* __restrict__ points, int n_frames,
    const int32_t* __restrict__ voxel_offsets, int64_t max_voxels, const int32_t* __restrict__ cnt,
    const int32_t* __restrict__ seg, const int32_t* __restrict__ pidx, float* __restrict__ feature,
    long long* __restrict__ coords, long long* __restrict__ counts) {
  __shared__ float4 stage[kFeatWarps][kMaxT];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float4* pts = stage[warp];
  const int T = g.T;
  int64_t total = voxel_offsets[n_frames];
  if (total > max_voxels) total = max_voxels;
  const int64_t nwarps = (int64_t)gridDim.x * kFeatWarps;
  for (int64_t row = (int64_t)blockIdx.x * kFeatWarps + warp; row < total; row += nwarps) {
    const int m = warp_select_points(points, pidx, cnt[row], seg[row], T, pts, lane);
    double mx, my, mz;
    warp_centroid(pts, m, T, lane, mx, my, mz);
    float* out = feature + row * (int64_t)(T * 7);
    for (int e = lane; e < T * 7; e += 32) {
      const int t = e / 7, j = e - t * 7;
      out[e] = feature_value(pts, m, t, j, mx, my, mz);
    }
    if (lane == 0) {
      write_coords(g, pts[0], n_frames, voxel_offsets, row, coords);
      counts[row] = m;
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------
// Fused P7 + stacked VFE, packed: the [K,T,7] feature buffer is never written or read, and a warp
// works on the rows of SEVERAL voxels at once (KITTI voxels hold 4.4 points on average, one voxel
// per warp leaves most lanes idle outside the VFE-2 loop).
//
// A warp owns 32 consecutive voxel rows (lane <-> voxel for counts, segment starts, coordinates) and
// cuts them into batches of whole voxels with <= 32 DISTINCT feature rows in total (a voxel with m <
// T points has m real rows plus one representative of its T - m identical pad rows: max() ignores
// multiplicity).  Inside a batch lane <-> row: bucket load, first-T order (rank of the
// point index among the voxel's indices), 7-d features with the fp64 centroid arithmetic, mask, VFE-1.
// agg1 is a (voxel, unit)-parallel pass over shared memory; VFE-2 runs voxel by voxel with lane <->
// two output units and the rows broadcast from shared memory.  Voxels with more than 31 points (or
// more than T) take the one-voxel-per-warp path (warp_select_points / svfe_eval_voxel).
// Bit-identical to k_vox_features followed by k_svfe_fused (same values through the same operations).
// ---------------------------------------------------------------------------------------------
constexpr int kPackWarps = 8;
constexpr int kPackRange = 8;   // voxels per work item (small: ~6 k voxels per frame over 2,368 resident warps)

struct Vfe1Packed {   // VFE-1 parameters laid out for 128-bit broadcast loads
  float4 wlo[16];     // w[u][0..3]
  float4 whi[16];     // w[u][4..6], bias
  float s[16], t[16];
};

struct PackWarp {
  float x[32][kXS];             // batch rows: VFE-1 output
  float4 pts[32];               // the batch's points in slot order (row r of a voxel = its r-th point)
  double mean[kPackRange][3];   // per batch voxel: centroid (utils.py:87-88)
  float agg[kPackRange][16];    // per batch voxel: agg1
  int rv[32];                   // row -> owning voxel (lane of the work item), start markers
  uint8_t first[kPackRange], nrows[kPackRange], npts[kPackRange];  // per batch voxel
};

__device__ __forceinline__ size_t pack_warp_bytes(int T) {
  const size_t rows_bytes = ((size_t)(T + 1) * 7 * 4 + 15) & ~(size_t)15;
  const size_t slow = (size_t)T * 16 + rows_bytes + (size_t)(T + 1) * kXS * 4;
  const size_t fast = sizeof(PackWarp);
  return ((slow > fast ? slow : fast) + 15) & ~(size_t)15;
}

__global__ void __launch_bounds__(kPackWarps * 32, 3) k_vox_svfe_packed(
    const vn_grid_t g, const float4* __restrict__ points, int n_frames,
    const int32_t* __restrict__ voxel_offsets, int64_t max_voxels, const int32_t* __restrict__ cnt,
    const int32_t* __restrict__ seg, const int32_t* __restrict__ pidx, const float* __restrict__ w1,
    const float* __restrict__ b1, const float* __restrict__ s1, const float* __restrict__ t1,
    const float* __restrict__ w2, const float* __restrict__ b2, const float* __restrict__ s2,
    const float* __restrict__ t2, float* __restrict__ voxelwise, __nv_bfloat16* __restrict__ dense,
    long long* __restrict__ coords, long long* __restrict__ counts, int32_t* __restrict__ counters,
    const int32_t* __restrict__ biglist, int dense_rows) {
  // dense_rows != 0: `dense` is the compact (K,128) bf16 row table (row = voxel row) instead of the (B,D,H,W,128) grid
  extern __shared__ __align__(16) uint8_t dyn_smem[];
  __shared__ Vfe1Params p1;   // one-voxel path
  __shared__ Vfe1Packed pk;   // packed path
  __shared__ Vfe2Shared sh;
  const int T = g.T;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint8_t* wbase = dyn_smem + warp * pack_warp_bytes(T);
  PackWarp& W = *reinterpret_cast<PackWarp*>(wbase);
  load_vfe1_params(p1, w1, b1, s1, t1);
  if (threadIdx.x < 16) {
    const int u = threadIdx.x;
    pk.wlo[u] = make_float4(w1[u * 7 + 0], w1[u * 7 + 1], w1[u * 7 + 2], w1[u * 7 + 3]);
    pk.whi[u] = make_float4(w1[u * 7 + 4], w1[u * 7 + 5], w1[u * 7 + 6], b1[u]);
    pk.s[u] = s1[u];
    pk.t[u] = t1[u];
  }
  Vfe2Lane q;
  load_vfe2(q, sh, w2, b2, s2, t2, lane);
  __syncthreads();

  int64_t total = voxel_offsets[n_frames];
  if (total > max_voxels) total = max_voxels;
  // ---- heavy items first: voxels with > 31 (or > T) points, one voxel per warp (they sit next to each
  // other in row order - dense cells near the sensor - so leaving them inside the 8-voxel items makes
  // stragglers) --------------------------------------------------------------------------------------
  {
    const int nbig = counters[kCtrBigCount];
    for (;;) {
      int bi = 0;
      if (lane == 0) bi = atomicAdd(&counters[kCtrBigNext], 1);
      bi = __shfl_sync(0xffffffffu, bi, 0);
      if (bi >= nbig) break;
      const int64_t vrow = biglist[bi];
      if (vrow >= total) continue;
      const int c_v = cnt[vrow], s_v = seg[vrow];
      const size_t rows_bytes = ((size_t)(T + 1) * 7 * 4 + 15) & ~(size_t)15;
      float4* pts = reinterpret_cast<float4*>(wbase);
      float* rows = reinterpret_cast<float*>(wbase + (size_t)T * 16);
      float* x1 = reinterpret_cast<float*>(wbase + (size_t)T * 16 + rows_bytes);
      const int m = warp_select_points(points, pidx, c_v, s_v, T, pts, lane, reinterpret_cast<int*>(rows),
                                       (int)((pack_warp_bytes(T) - (size_t)T * 16) / 4));
      __syncwarp();
      int64_t cell_v = 0;
      if (lane == 0) {
        cell_v = write_coords(g, pts[0], n_frames, voxel_offsets, vrow, coords);
        if (counts != nullptr) counts[vrow] = m;
      }
      cell_v = __shfl_sync(0xffffffffu, cell_v, 0);
      double mx, my, mz;
      warp_centroid(pts, m, T, lane, mx, my, mz);
      __syncwarp();
      const int Rv = m < T ? m + 1 : T;  // real rows 0..m-1, then one pad row at index m
      for (int e = lane; e < Rv * 7; e += 32) {
        const int t = e / 7, jx = e - t * 7;
        rows[e] = feature_value(pts, m, t, jx, mx, my, mz);
      }
      __syncwarp();
      float o[4];
      svfe_eval_voxel(rows, m, Rv, m, x1, p1, q, sh, o, lane);
      store_voxel_out(o, lane, voxelwise != nullptr ? voxelwise + vrow * 128 : nullptr,
                      dense != nullptr ? dense + (dense_rows ? vrow : cell_v) * 128 : nullptr);
      __syncwarp();
    }
  }
  const int64_t nranges = (total + kPackRange - 1) / kPackRange;
  // work items are pulled from a counter; the next item is requested before the current one is processed
  int32_t* work_counter = counters + kCtrItems;
  int nxt = 0;
  if (lane == 0) nxt = atomicAdd(work_counter, 1);
  for (;;) {
    const int64_t rg = __shfl_sync(0xffffffffu, nxt, 0);
    if (rg >= nranges) break;
    if (lane == 0) nxt = atomicAdd(work_counter, 1);
    // ---- lane <-> voxel: counts, bucket start, coordinates ----------------------------------------
    const int64_t row = rg * kPackRange + lane;
    const bool valid = (int)lane < kPackRange && row < total;
    int c = 0, s0 = 0;
    int64_t cell = 0;
    if (valid) {
      c = cnt[row];
      s0 = seg[row];
    }
    const bool big = valid && (c > 31 || c > T);  // done above
    if (valid && !big) {
      cell = write_coords(g, points[pidx[s0]], n_frames, voxel_offsets, row, coords);
      if (counts != nullptr) counts[row] = c;
    }
    const int R = (valid && !big) ? c + (c < T ? 1 : 0) : 0;  // distinct rows of this voxel (<= 32)
    const int incl = (int)warp_incl_scan((uint32_t)R, lane);
    const int excl = incl - R;
    const int Rall = __shfl_sync(0xffffffffu, incl, 31);
    int base = 0, a = 0;
    while (base < Rall) {
      // ---- batch = voxels [a, b_end): as many whole voxels as fit 32 rows -------------------------
      const int b_end = min(__popc(__ballot_sync(0xffffffffu, incl - base <= 32)), kPackRange);
      const int Rtot = __shfl_sync(0xffffffffu, incl, b_end - 1) - base;
      const int nvox = b_end - a;
      W.rv[lane] = -1;
      __syncwarp();
      const bool owner = (int)lane >= a && (int)lane < b_end;
      if (owner) {
        if (R > 0) W.rv[excl - base] = (int)lane;
        W.first[lane - a] = (uint8_t)(excl - base);
        W.nrows[lane - a] = (uint8_t)R;
        W.npts[lane - a] = (uint8_t)c;
      }
      __syncwarp();
      // ---- lane <-> row --------------------------------------------------------------------------
      int w = W.rv[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {  // owners are ascending: inclusive max-scan fills the gaps
        const int t = __shfl_up_sync(0xffffffffu, w, d);
        if ((int)lane >= d) w = max(w, t);
      }
      const bool rowvalid = (int)lane < Rtot;
      const int wsrc = rowvalid ? w : 0;
      const int c_w = __shfl_sync(0xffffffffu, c, wsrc);
      const int s_w = __shfl_sync(0xffffffffu, s0, wsrc);
      const int first_w = __shfl_sync(0xffffffffu, excl, wsrc) - base;
      const int j = (int)lane - first_w;          // row index inside the voxel; j == c_w is the pad row
      const bool real = rowvalid && j < c_w;
      int idx = 0x7fffffff;
      float4 p = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      if (real) {
        idx = pidx[s_w + j];
        p = points[idx];
      }
      // first-T rule (utils.py:78-84): slot = rank of the point index among the voxel's indices
      const int maxc = __reduce_max_sync(0xffffffffu, real ? c_w : 0);
      int rank = 0;
      for (int jj = 0; jj < maxc; ++jj) {
        const int o = __shfl_sync(0xffffffffu, idx, (first_w + jj) & 31);
        rank += (jj < c_w && o < idx) ? 1 : 0;
      }
      if (real) W.pts[first_w + rank] = p;
      __syncwarp();
      // centroid (utils.py:87-88): fp32 sum in slot order incl. the +0.0 of the first pad, fp64 divide
      for (int e0 = 0; e0 < nvox * 3; e0 += 32) {
        const int e = e0 + (int)lane;
        if (e < nvox * 3) {
          const int k = e / 3, comp = e - 3 * k;
          const int n = W.npts[k];
          if (W.nrows[k] != 0) {
            const float* src = reinterpret_cast<const float*>(&W.pts[W.first[k]]) + comp;
            float sum = src[0];
            for (int t = 1; t < n; ++t) sum = __fadd_rn(sum, src[4 * t]);
            if (n < T) sum = __fadd_rn(sum, 0.0f);
            W.mean[k][comp] = (double)sum / (double)n;
          }
        }
      }
      __syncwarp();
      // 7-d features of my row (utils.py:71, 87-88), mask (model.py:95-96), VFE-1 (model.py:74-76)
      bool on = false;
      if (rowvalid) {
        const int k = w - a;
        const float4 pt = real ? W.pts[lane] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        float f[7];
        f[0] = pt.x; f[1] = pt.y; f[2] = pt.z; f[3] = pt.w;
        f[4] = __double2float_rn((double)pt.x - W.mean[k][0]);
        f[5] = __double2float_rn((double)pt.y - W.mean[k][1]);
        f[6] = __double2float_rn((double)pt.z - W.mean[k][2]);
        float mx = f[0];
#pragma unroll
        for (int i = 1; i < 7; ++i) mx = fmaxf(mx, f[i]);
        on = mx != 0.0f;
        float4* xrow = reinterpret_cast<float4*>(W.x[lane]);
#pragma unroll
        for (int u4 = 0; u4 < 4; ++u4) {
          float v[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int u = u4 * 4 + i;
            const float4 lo = pk.wlo[u], hi = pk.whi[u];
            float acc = hi.w;
            acc = fmaf(lo.x, f[0], acc);
            acc = fmaf(lo.y, f[1], acc);
            acc = fmaf(lo.z, f[2], acc);
            acc = fmaf(lo.w, f[3], acc);
            acc = fmaf(hi.x, f[4], acc);
            acc = fmaf(hi.y, f[5], acc);
            acc = fmaf(hi.z, f[6], acc);
            v[i] = fmaf(fmaxf(acc, 0.0f), pk.s[u], pk.t[u]);
          }
          xrow[u4] = make_float4(v[0], v[1], v[2], v[3]);
        }
      }
      const uint32_t on_bits = __ballot_sync(0xffffffffu, on);
      __syncwarp();
      // agg1[k][u] = max over the voxel's rows (all T rows of the reference: pads are represented)
      for (int e = lane; e < nvox * 16; e += 32) {
        const int k = e >> 4, u = e & 15;
        const int n = W.nrows[k];
        if (n != 0) {
          const int f0 = W.first[k];
          float m = W.x[f0][u];
          for (int t = 1; t < n; ++t) m = fmaxf(m, W.x[f0 + t][u]);
          W.agg[k][u] = m;
        }
      }
      __syncwarp();
      // ---- VFE-2 + voxel-wise max, voxel by voxel (lane <-> output units) -----------------------------
      for (int k = 0; k < nvox; ++k) {
        const int n = W.nrows[k];
        const int64_t cell_k = __shfl_sync(0xffffffffu, cell, a + k);
        if (n == 0) continue;
        const int f0 = W.first[k];
        const uint32_t bits = (on_bits >> f0) & (n >= 32 ? 0xffffffffu : ((1u << n) - 1u));
        float o[4];
        vfe2_voxel_max<uint32_t>(W.x[f0], W.agg[k], n, bits, q, sh, lane, o);  // (a batch holds at most 32 rows)
        const int64_t vrow = rg * kPackRange + a + k;
        store_voxel_out(o, lane, voxelwise != nullptr ? voxelwise + vrow * 128 : nullptr,
                        dense != nullptr ? dense + (dense_rows ? vrow : cell_k) * 128 : nullptr);
      }
      __syncwarp();
      base += Rtot;
      a = b_end;
    }
  }
}

int check_common(const vn_grid_t* grid, const void* points, const int32_t* frame_offsets, int32_t n_frames,
                 int64_t n_points_total, const void* workspace, size_t workspace_bytes) {
  VN_REQUIRE(grid != nullptr, "voxelize: grid is null");
  VN_REQUIRE(grid->D > 0 && grid->H > 0 && grid->W > 0, "voxelize: empty grid");
  VN_REQUIRE((int64_t)grid->D * grid->H * grid->W < (int64_t)0x7fffffff, "voxelize: grid too large");
  VN_REQUIRE(grid->T >= 1 && grid->T <= kMaxT, "voxelize: T=%d outside [1,%d]", grid->T, kMaxT);
  VN_REQUIRE(n_frames >= 1, "voxelize: n_frames=%d", n_frames);
  VN_REQUIRE(n_points_total >= 0 && n_points_total < (int64_t)0x7fffffff, "voxelize: bad point count");
  VN_REQUIRE(frame_offsets != nullptr, "voxelize: frame_offsets is null");
  VN_REQUIRE(points != nullptr || n_points_total == 0, "voxelize: points is null");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(points) & 15) == 0, "voxelize: points must be 16-byte aligned");
  VN_REQUIRE(workspace != nullptr, "voxelize: workspace is null");
  const size_t need = carve(*grid, n_frames, n_points_total, nullptr).total_bytes;
  if (workspace_bytes < need) {
    set_error("voxelize: workspace %zu < %zu bytes", workspace_bytes, need);
    return VN_ERR_WORKSPACE;
  }
  VN_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "voxelize: workspace must be 256-byte aligned");
  return VN_OK;
}

// Coordinates of every voxel row right after the bucketing (one thread per row: the voxel's cell is the cell of
// the first point of its bucket).  The fused voxel->SVFE kernel writes the same values again later; having them
// early lets everything that depends on the coordinates only (live-box marking, conv planning) run next to it.
__global__ void __launch_bounds__(256) k_vox_coords(const vn_grid_t g, const float4* __restrict__ points, int n_frames,
                                                   const int32_t* __restrict__ voxel_offsets, int64_t max_voxels,
                                                   const int32_t* __restrict__ seg, const int32_t* __restrict__ pidx,
                                                   long long* __restrict__ coords) {
  int64_t total = voxel_offsets[n_frames];
  if (total > max_voxels) total = max_voxels;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; row < total; row += stride)
    write_coords(g, points[pidx[seg[row]]], n_frames, voxel_offsets, row, coords);
}

// P5 + P6: CSR segments per voxel and the bucketed point indices.
int launch_bucket(const VoxWorkspace& w, const int32_t* frame_offsets, int n_frames, int64_t n_points_total,
                  const int32_t* voxel_offsets, int T, cudaStream_t st) {
  VN_CUDA(cudaMemsetAsync(w.counters, 0, 64 * sizeof(int32_t), st));
  const int seg_blocks = (int)(w.seg_chunks < 64 ? w.seg_chunks : 64);
  k_vox_seg_partials<<<dim3(seg_blocks, n_frames), 256, 0, st>>>(voxel_offsets, w.cnt, (int)w.seg_chunks, w.segpart);
  VN_LAUNCH_CHECK("k_vox_seg_partials");
  k_vox_segment_scan<<<dim3(seg_blocks, n_frames), 256, 0, st>>>(frame_offsets, voxel_offsets, w.cnt, (int)w.seg_chunks, w.segpart,
                                                                 w.seg, w.counters, w.biglist, T);
  VN_LAUNCH_CHECK("k_vox_segment_scan");
  const int64_t want = (n_points_total + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 16;
  k_vox_place<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(n_points_total, w.keyrow, w.arrival, w.seg, w.pidx);
  VN_LAUNCH_CHECK("k_vox_place");
  return VN_OK;
}

}  // namespace
}  // namespace vn

using namespace vn;

namespace vn {
int occupancy_views(const vn_grid_t& grid, int n_frames, int64_t n_points_total, const void* workspace, size_t workspace_bytes,
                    const uint32_t** bitmap, const uint32_t** wordprefix, size_t* words) {
  const VoxWorkspace w = carve(grid, n_frames, n_points_total, const_cast<void*>(workspace));
  if (workspace_bytes < w.total_bytes) return -1;
  *bitmap = w.bitmap;
  *wordprefix = w.wordprefix;
  *words = w.words;
  return 0;
}
}  // namespace vn

extern "C" {

size_t vn_voxelize_workspace_bytes(const vn_grid_t* grid, int32_t n_frames, int64_t n_points_total) {
  if (grid == nullptr || n_frames < 1 || n_points_total < 0) return 0;
  return carve(*grid, n_frames, n_points_total, nullptr).total_bytes;
}

int vn_voxelize_index(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets,
                      int32_t n_frames, int64_t n_points_total, int32_t max_points_per_frame,
                      int32_t* voxel_offsets, void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  int rc = check_common(grid, points, frame_offsets, n_frames, n_points_total, workspace, workspace_bytes);
  if (rc != VN_OK) return rc;
  VN_REQUIRE(voxel_offsets != nullptr, "voxelize_index: voxel_offsets is null");
  VN_REQUIRE(max_points_per_frame >= 0, "voxelize_index: max_points_per_frame < 0");
  cudaStream_t st = as_stream(stream);
  const VoxWorkspace w = carve(*grid, n_frames, n_points_total, workspace);
  const int sms = sm_count();

  {
    const size_t n16 = w.zero_bytes / 16;
    const int blocks = (int)((n16 + 255) / 256 < (size_t)sms * 8 ? (n16 + 255) / 256 : (size_t)sms * 8);
    k_fill_zero16<<<blocks, 256, 0, st>>>(reinterpret_cast<uint4*>(workspace), n16);
    VN_LAUNCH_CHECK("k_fill_zero16");
  }
  const int bx = max_points_per_frame > 0 ? (max_points_per_frame + 255) / 256 : 1;
  if (n_points_total > 0) {
    k_vox_mark<<<dim3(bx, n_frames), 256, 0, st>>>(*grid, reinterpret_cast<const float4*>(points),
                                                   frame_offsets, w.bitmap, w.words, w.keyrow);
    VN_LAUNCH_CHECK("k_vox_mark");
  }
  k_vox_word_partials<<<dim3((unsigned)w.word_chunks, n_frames), 256, 0, st>>>(w.bitmap, w.words, (int)w.word_chunks, w.wordpart);
  VN_LAUNCH_CHECK("k_vox_word_partials");
  k_vox_word_scan<<<dim3((unsigned)w.word_chunks, n_frames), 256, 0, st>>>(w.bitmap, w.words, (int)w.word_chunks, w.wordpart,
                                                                            w.wordprefix, w.frame_k);
  VN_LAUNCH_CHECK("k_vox_word_scan");
  k_vox_frame_offsets<<<1, 32, 0, st>>>(w.frame_k, n_frames, voxel_offsets);
  VN_LAUNCH_CHECK("k_vox_frame_offsets");
  if (n_points_total > 0) {
    k_vox_count<<<dim3(bx, n_frames), 256, 0, st>>>(frame_offsets, w.bitmap, w.wordprefix, w.words,
                                                    voxel_offsets, w.keyrow, w.arrival, w.cnt);
    VN_LAUNCH_CHECK("k_vox_count");
  }
  return VN_OK;
}

int vn_voxelize_fill(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets,
                     int32_t n_frames, int64_t n_points_total, int32_t max_points_per_frame,
                     const int32_t* voxel_offsets, int64_t max_voxels, float* feature_buffer,
                     int64_t* coordinate_buffer, int64_t* number_buffer, void* workspace,
                     size_t workspace_bytes, vn_stream_t stream) {
  int rc = check_common(grid, points, frame_offsets, n_frames, n_points_total, workspace, workspace_bytes);
  if (rc != VN_OK) return rc;
  (void)max_points_per_frame;
  VN_REQUIRE(voxel_offsets != nullptr, "voxelize_fill: voxel_offsets is null");
  VN_REQUIRE(max_voxels >= 0, "voxelize_fill: max_voxels < 0");
  if (max_voxels == 0 || n_points_total == 0) return VN_OK;
  VN_REQUIRE(feature_buffer && coordinate_buffer && number_buffer, "voxelize_fill: null output");
  cudaStream_t st = as_stream(stream);
  const VoxWorkspace w = carve(*grid, n_frames, n_points_total, workspace);
  const int sms = sm_count();
  rc = launch_bucket(w, frame_offsets, n_frames, n_points_total, voxel_offsets, grid->T, st);
  if (rc != VN_OK) return rc;
  {
    // persistent warps: 8 resident blocks of 8 warps per SM
    const int64_t want = (max_voxels + kFeatWarps - 1) / kFeatWarps;
    const int blocks = (int)(want < (int64_t)sms * 8 ? want : (int64_t)sms * 8);
    k_vox_features<<<blocks, kFeatWarps * 32, 0, st>>>(
        *grid, reinterpret_cast<const float4*>(points), n_frames, voxel_offsets, max_voxels, w.cnt, w.seg,
        w.pidx, feature_buffer, reinterpret_cast<long long*>(coordinate_buffer),
        reinterpret_cast<long long*>(number_buffer));
    VN_LAUNCH_CHECK("k_vox_features");
  }
  return VN_OK;
}

static int svfe_fused_impl(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets, int32_t n_frames,
                           int64_t n_points_total, const int32_t* voxel_offsets, int64_t max_voxels, const float* w1,
                           const float* b1, const float* s1, const float* t1, const float* w2, const float* b2,
                           const float* s2, const float* t2, float* voxelwise, void* dense_bf16,
                           int64_t* coordinate_buffer, int64_t* number_buffer, void* workspace, size_t workspace_bytes,
                           vn_stream_t stream, bool bucketed, bool dense_rows = false) {
  int rc = check_common(grid, points, frame_offsets, n_frames, n_points_total, workspace, workspace_bytes);
  if (rc != VN_OK) return rc;
  VN_REQUIRE(voxel_offsets != nullptr, "voxelize_svfe_fused: voxel_offsets is null");
  VN_REQUIRE(max_voxels >= 0, "voxelize_svfe_fused: max_voxels < 0");
  if (max_voxels == 0 || n_points_total == 0) return VN_OK;
  VN_REQUIRE(w1 && b1 && s1 && t1 && w2 && b2 && s2 && t2, "voxelize_svfe_fused: null parameter pointer");
  VN_REQUIRE((voxelwise || dense_bf16) && coordinate_buffer, "voxelize_svfe_fused: null output");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(dense_bf16) & 3) == 0 && (reinterpret_cast<uintptr_t>(coordinate_buffer) & 15) == 0,
             "voxelize_svfe_fused: misaligned output");
  cudaStream_t st = as_stream(stream);
  const VoxWorkspace w = carve(*grid, n_frames, n_points_total, workspace);
  if (!bucketed) {  // else vn_voxelize_bucket_coords ran: segments, buckets and the zeroed work counters are in place
    rc = launch_bucket(w, frame_offsets, n_frames, n_points_total, voxel_offsets, grid->T, st);
    if (rc != VN_OK) return rc;
  }
  const int T = grid->T;
  const size_t rows_bytes = ((size_t)(T + 1) * 7 * 4 + 15) & ~(size_t)15;
  const size_t slow = (size_t)T * 16 + rows_bytes + (size_t)(T + 1) * kXS * 4;
  const size_t per_warp = ((slow > sizeof(PackWarp) ? slow : sizeof(PackWarp)) + 15) & ~(size_t)15;
  const size_t smem = kPackWarps * per_warp;
  VN_CUDA(cudaFuncSetAttribute(k_vox_svfe_packed, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t want = (max_voxels + kPackRange * kPackWarps - 1) / (kPackRange * kPackWarps);
  const int64_t cap = (int64_t)sm_count() * 3;
  k_vox_svfe_packed<<<(unsigned)(want < cap ? want : cap), kPackWarps * 32, smem, st>>>(
      *grid, reinterpret_cast<const float4*>(points), n_frames, voxel_offsets, max_voxels, w.cnt, w.seg, w.pidx, w1, b1,
      s1, t1, w2, b2, s2, t2, voxelwise, reinterpret_cast<__nv_bfloat16*>(dense_bf16),
      bucketed ? nullptr : reinterpret_cast<long long*>(coordinate_buffer), reinterpret_cast<long long*>(number_buffer),
      w.counters, w.biglist, dense_rows ? 1 : 0);
  VN_LAUNCH_CHECK("k_vox_svfe_packed");
  return VN_OK;
}

int vn_voxelize_svfe_fused(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets, int32_t n_frames,
                           int64_t n_points_total, const int32_t* voxel_offsets, int64_t max_voxels, const float* w1,
                           const float* b1, const float* s1, const float* t1, const float* w2, const float* b2,
                           const float* s2, const float* t2, float* voxelwise, void* dense_bf16,
                           int64_t* coordinate_buffer, int64_t* number_buffer, void* workspace, size_t workspace_bytes,
                           vn_stream_t stream) {
  return svfe_fused_impl(grid, points, frame_offsets, n_frames, n_points_total, voxel_offsets, max_voxels, w1, b1, s1, t1, w2,
                         b2, s2, t2, voxelwise, dense_bf16, coordinate_buffer, number_buffer, workspace, workspace_bytes,
                         stream, false);
}

int vn_voxelize_svfe_bucketed(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets, int32_t n_frames,
                              int64_t n_points_total, const int32_t* voxel_offsets, int64_t max_voxels, const float* w1,
                              const float* b1, const float* s1, const float* t1, const float* w2, const float* b2,
                              const float* s2, const float* t2, float* voxelwise, void* dense_bf16,
                              int64_t* coordinate_buffer, int64_t* number_buffer, void* workspace,
                              size_t workspace_bytes, vn_stream_t stream) {
  return svfe_fused_impl(grid, points, frame_offsets, n_frames, n_points_total, voxel_offsets, max_voxels, w1, b1, s1, t1, w2,
                         b2, s2, t2, voxelwise, dense_bf16, coordinate_buffer, number_buffer, workspace, workspace_bytes,
                         stream, true);
}

int vn_voxelize_svfe_rows(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets, int32_t n_frames,
                          int64_t n_points_total, const int32_t* voxel_offsets, int64_t max_voxels, const float* w1,
                          const float* b1, const float* s1, const float* t1, const float* w2, const float* b2,
                          const float* s2, const float* t2, void* rows_bf16, int64_t* coordinate_buffer, int32_t bucketed,
                          void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  VN_REQUIRE(rows_bf16 != nullptr && (reinterpret_cast<uintptr_t>(rows_bf16) & 15) == 0, "voxelize_svfe_rows: rows null or misaligned");
  return svfe_fused_impl(grid, points, frame_offsets, n_frames, n_points_total, voxel_offsets, max_voxels, w1, b1, s1, t1, w2,
                         b2, s2, t2, nullptr, rows_bf16, coordinate_buffer, nullptr, workspace, workspace_bytes, stream,
                         bucketed != 0, true);
}

int vn_voxelize_bucket_coords(const vn_grid_t* grid, const float* points, const int32_t* frame_offsets, int32_t n_frames,
                              int64_t n_points_total, const int32_t* voxel_offsets, int64_t max_voxels,
                              int64_t* coordinate_buffer, void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  int rc = check_common(grid, points, frame_offsets, n_frames, n_points_total, workspace, workspace_bytes);
  if (rc != VN_OK) return rc;
  VN_REQUIRE(voxel_offsets != nullptr && max_voxels >= 0, "voxelize_bucket_coords: bad argument");
  if (max_voxels == 0 || n_points_total == 0) return VN_OK;
  VN_REQUIRE(coordinate_buffer != nullptr && (reinterpret_cast<uintptr_t>(coordinate_buffer) & 15) == 0,
             "voxelize_bucket_coords: coordinate_buffer null or misaligned");
  cudaStream_t st = as_stream(stream);
  const VoxWorkspace w = carve(*grid, n_frames, n_points_total, workspace);
  rc = launch_bucket(w, frame_offsets, n_frames, n_points_total, voxel_offsets, grid->T, st);
  if (rc != VN_OK) return rc;
  const int64_t want = (max_voxels + 255) / 256;
  const int64_t cap = (int64_t)sm_count() * 8;
  k_vox_coords<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(*grid, reinterpret_cast<const float4*>(points), n_frames,
                                                                  voxel_offsets, max_voxels, w.seg, w.pidx,
                                                                  reinterpret_cast<long long*>(coordinate_buffer));
  VN_LAUNCH_CHECK("k_vox_coords");
  return VN_OK;
}

int vn_voxelize_counts_host(const int32_t* voxel_offsets, int32_t n_frames, int32_t* voxel_offsets_host,
                            vn_stream_t stream) {
  VN_REQUIRE(voxel_offsets && voxel_offsets_host && n_frames >= 1, "voxelize_counts_host: bad argument");
  cudaStream_t st = as_stream(stream);
  VN_CUDA(cudaMemcpyAsync(voxel_offsets_host, voxel_offsets, sizeof(int32_t) * (size_t)(n_frames + 1),
                          cudaMemcpyDeviceToHost, st));
  VN_CUDA(cudaStreamSynchronize(st));
  return VN_OK;
}

int vn_fill_zero(void* ptr, size_t bytes, vn_stream_t stream) {
  VN_REQUIRE(ptr != nullptr || bytes == 0, "fill_zero: null pointer");
  VN_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (bytes & 15) == 0,
             "fill_zero: pointer and size must be multiples of 16");
  if (bytes == 0) return VN_OK;
  const size_t n16 = bytes / 16;
  const size_t want = (n16 + 255) / 256;
  const size_t cap = (size_t)sm_count() * 16;
  k_fill_zero16<<<(unsigned)(want < cap ? want : cap), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<uint4*>(ptr), n16);
  VN_LAUNCH_CHECK("k_fill_zero16");
  return VN_OK;
}

}  // extern "C"
