// RPN target assignment on the device: the anchor x ground-truth IoU matrix and the positive / negative /
// regression targets of one batch.  Replaces the pure-Python double loop `bbox_iou` (reference
// voxelnet/utils.py:344-373, N*K interpreter iterations per frame, N = 70,400 anchors for Car) and the per-frame body of
// `generate_targets` (utils.py:395-462).
//
// Parity rules (tests/test_gpu_targets.py):
// * the IoU is float32 arithmetic in the reference's operation order - explicit round-to-nearest intrinsics, so the
//   compiler cannot contract a*b+c into an FMA - and is bit-identical, including the reference's anchor-area term
//   (box1[n,1] - box1[n,0] + 1) * (box1[n,3] - box1[n,1] + 1) (utils.py:366-367) as written there;
// * positives / negatives are therefore identical index sets; the regression targets are float64 like the
//   reference's arrays: differences, quotients and the sqrt are IEEE-exact, the three log() terms agree to 1 ulp.
//
// Selection rules restated (utils.py:411-430): an anchor is positive if its IoU with any box exceeds pos_iou, or if it
// is the arg-max anchor (lowest index among equals, np.argmax) of a box whose best IoU is > 0.  Its box is the
// lowest-numbered box above the threshold, else the lowest-numbered box it is the arg-max of (np.unique(...,
// return_index=True) keeps the first occurrence of an anchor in [where(iou > pos) ..., argmax ...]).  An anchor is
// negative if its IoU with every box is below neg_iou (all anchors when the frame has no box).  Positives are not
// removed from the negatives (the reference's code for that is commented out, utils.py:453-460).
#include "vn_common.cuh"

#include <mutex>

// Scratch of vn_rpn_targets_static: lives in every CUDA context that loads the library, from load until exit.
// In host code `g_targets_scratch` is only a stand-in for the symbol - never a device address: the address comes from
// cudaGetSymbolAddress, once per device, with that device current (static_scratch_for_current_device below).
static __device__ __align__(256) unsigned char g_targets_scratch[VN_RPN_TARGETS_STATIC_BYTES];

namespace vn {
namespace {

struct F4 { float x1, y1, x2, y2; };

// One IoU entry, utils.py:352-371 in float32.
__device__ __forceinline__ float iou_entry(const F4& a, const F4& g, float g_area) {
  const float iw = __fadd_rn(__fsub_rn(fminf(a.x2, g.x2), fmaxf(a.x1, g.x1)), 1.0f);
  if (!(iw > 0.0f)) return 0.0f;
  const float ih = __fadd_rn(__fsub_rn(fminf(a.y2, g.y2), fmaxf(a.y1, g.y1)), 1.0f);
  if (!(ih > 0.0f)) return 0.0f;
  const float a_term = __fmul_rn(__fadd_rn(__fsub_rn(a.y1, a.x1), 1.0f), __fadd_rn(__fsub_rn(a.y2, a.y1), 1.0f));
  const float inter = __fmul_rn(iw, ih);
  const float ua = __fsub_rn(__fadd_rn(a_term, g_area), inter);
  return __fdiv_rn(inter, ua);
}

__device__ __forceinline__ float box_area(const F4& g) {
  return __fmul_rn(__fadd_rn(__fsub_rn(g.x2, g.x1), 1.0f), __fadd_rn(__fsub_rn(g.y2, g.y1), 1.0f));
}

__global__ void __launch_bounds__(256)
    k_bbox_iou(const F4* __restrict__ box1, const F4* __restrict__ box2, int64_t N, int K, float* __restrict__ out) {
  const int64_t total = N * K;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / K;
    const int k = (int)(i - n * K);
    const F4 g = box2[k];
    out[i] = iou_entry(box1[n], g, box_area(g));
  }
}

constexpr int kMaxBoxesSmem = 1024;  // boxes of one frame staged per pass

// Pass 1: per anchor, the first box above pos_iou and the all-below-neg flag; per box, the arg-max anchor as one
// 64-bit atomicMax key (IoU bits high, inverted anchor index low: highest IoU, then lowest index).
__global__ void __launch_bounds__(256)
    k_targets_scan(const F4* __restrict__ anchors_standup, int A, const F4* __restrict__ gt_standup,
                   const int32_t* __restrict__ gt_offsets, float pos_iou, float neg_iou, int32_t* __restrict__ pos_box,
                   double* __restrict__ neg, unsigned long long* __restrict__ best) {
  __shared__ F4 s_box[kMaxBoxesSmem];
  __shared__ float s_area[kMaxBoxesSmem];
  __shared__ unsigned long long s_best[kMaxBoxesSmem];
  const int b = blockIdx.y;
  const int g0 = gt_offsets[b], G = gt_offsets[b + 1] - g0;
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  F4 an = {0.f, 0.f, 0.f, 0.f};
  if (a < A) an = anchors_standup[a];
  int first_pos = -1;
  bool all_neg = true;
  for (int base = 0; base < G; base += kMaxBoxesSmem) {
    const int n = min(G - base, kMaxBoxesSmem);
    __syncthreads();
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
      const F4 g = gt_standup[g0 + base + j];
      s_box[j] = g;
      s_area[j] = box_area(g);
      s_best[j] = 0ull;
    }
    __syncthreads();
    if (a < A) {
      for (int j = 0; j < n; ++j) {
        const float v = iou_entry(an, s_box[j], s_area[j]);
        if (v > pos_iou && first_pos < 0) first_pos = base + j;
        if (!(v < neg_iou)) all_neg = false;
        if (v > 0.0f)
          atomicMax(&s_best[j], ((unsigned long long)__float_as_uint(v) << 32) | (0xFFFFFFFFu - (unsigned)a));
      }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < n; j += blockDim.x)
      if (s_best[j] != 0ull) atomicMax(&best[g0 + base + j], s_best[j]);
  }
  if (a < A) {
    pos_box[(int64_t)b * A + a] = first_pos;
    neg[(int64_t)b * A + a] = all_neg ? 1.0 : 0.0;
  }
}

// Pass 2: arg-max anchors claim their (lowest-numbered) box unless they already have one above the threshold.
__global__ void __launch_bounds__(256)
    k_targets_claim(const int32_t* __restrict__ gt_offsets, int n_frames, int A,
                    const unsigned long long* __restrict__ best, int32_t* __restrict__ claim) {
  const int total = gt_offsets[n_frames];
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < total; g += gridDim.x * blockDim.x) {
    const unsigned long long key = best[g];
    if (key == 0ull) continue;  // best IoU is 0: no arg-max positive (utils.py:413-414)
    int b = 0;
    while (gt_offsets[b + 1] <= g) ++b;
    const int a = (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
    atomicMin(&claim[(int64_t)b * A + a], g - gt_offsets[b]);
  }
}

// Pass 3: every output element is written (zeros for non-positives), so the caller does not clear anything.
// anchors [A,7] and gt_boxes [G,7] are (x, y, z, h, w, l, r) float64, utils.py:434-462.
__global__ void __launch_bounds__(256)
    k_targets_write(const double* __restrict__ anchors, int A, const double* __restrict__ gt_boxes,
                    const int32_t* __restrict__ gt_offsets, double anchor_h, const int32_t* __restrict__ pos_box,
                    const int32_t* __restrict__ claim, double* __restrict__ pos, double* __restrict__ targets) {
  const int b = blockIdx.y;
  const int a = blockIdx.x * blockDim.x + threadIdx.x;
  if (a >= A) return;
  const int64_t i = (int64_t)b * A + a;
  int g = pos_box[i];
  if (g < 0) {
    const int c = claim[i];
    g = c == INT32_MAX ? -1 : c;
  }
  double t[7] = {0, 0, 0, 0, 0, 0, 0};
  if (g >= 0) {
    const double* an = anchors + (int64_t)a * 7;
    const double* gt = gt_boxes + (int64_t)(gt_offsets[b] + g) * 7;
    const double diag = sqrt(__dadd_rn(__dmul_rn(an[4], an[4]), __dmul_rn(an[5], an[5])));
    t[0] = __ddiv_rn(__dsub_rn(gt[0], an[0]), diag);
    t[1] = __ddiv_rn(__dsub_rn(gt[1], an[1]), diag);
    t[2] = __ddiv_rn(__dsub_rn(gt[2], an[2]), anchor_h);
    t[3] = log(__ddiv_rn(gt[3], an[3]));
    t[4] = log(__ddiv_rn(gt[4], an[4]));
    t[5] = log(__ddiv_rn(gt[5], an[5]));
    t[6] = __dsub_rn(gt[6], an[6]);
  }
  pos[i] = g >= 0 ? 1.0 : 0.0;
#pragma unroll
  for (int j = 0; j < 7; ++j) targets[i * 7 + j] = t[j];
}

__global__ void __launch_bounds__(256) k_targets_init(unsigned long long* best, int64_t n_best, int32_t* claim, int64_t n_claim) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_claim; i += stride) claim[i] = INT32_MAX;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_best; i += stride) best[i] = 0ull;
}

constexpr size_t kStaticScratchBytes = VN_RPN_TARGETS_STATIC_BYTES;

constexpr int kMaxDevices = 64;
std::mutex g_scratch_mutex;
void* g_scratch_addr[kMaxDevices] = {};

// 0 and *out = the device address of g_targets_scratch on the current device, else a cudaError_t (> 0) / VN_ERR_*.
int static_scratch_for_current_device(void** out) {
  int dev = -1;
  VN_CUDA(cudaGetDevice(&dev));
  VN_REQUIRE(dev >= 0 && dev < kMaxDevices, "rpn_targets_static: device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (g_scratch_addr[dev] == nullptr) {
    void* p = nullptr;
    VN_CUDA(cudaGetSymbolAddress(&p, g_targets_scratch));
    VN_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "rpn_targets_static: bad symbol address");
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return VN_OK;
}

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

int vn_bbox_iou(const float* box1, const float* box2, int64_t N, int32_t K, float* overlaps, vn_stream_t stream) {
  VN_REQUIRE(N >= 0 && K >= 0, "bbox_iou: negative size");
  if (N == 0 || K == 0) return VN_OK;
  VN_REQUIRE(box1 && box2 && overlaps, "bbox_iou: null pointer");
  VN_REQUIRE(((reinterpret_cast<uintptr_t>(box1) | reinterpret_cast<uintptr_t>(box2)) & 15) == 0,
             "bbox_iou: boxes must be 16-byte aligned");
  const int64_t want = (N * K + 255) / 256, cap = (int64_t)sm_count() * 8;
  k_bbox_iou<<<(unsigned)(want < cap ? want : cap), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<const F4*>(box1), reinterpret_cast<const F4*>(box2), N, K, overlaps);
  VN_LAUNCH_CHECK("k_bbox_iou");
  return VN_OK;
}

size_t vn_rpn_targets_workspace_bytes(int32_t n_frames, int32_t n_anchors, int32_t n_boxes_total) {
  if (n_frames < 0 || n_anchors < 0 || n_boxes_total < 0) return 0;
  return align_up((size_t)(n_boxes_total > 0 ? n_boxes_total : 1) * 8, 16) + (size_t)n_frames * n_anchors * 4 * 2 + 16;
}

int vn_rpn_targets(const float* anchors_standup, const double* anchors, int32_t n_anchors, const float* gt_standup,
                   const double* gt_boxes, const int32_t* gt_offsets, int32_t n_frames, int32_t n_boxes_total,
                   float pos_iou, float neg_iou, double anchor_h, double* pos_equal_one, double* neg_equal_one,
                   double* targets, void* workspace, size_t workspace_bytes, vn_stream_t stream) {
  VN_REQUIRE(n_frames >= 0 && n_anchors >= 0 && n_boxes_total >= 0, "rpn_targets: negative size");
  if (n_frames == 0 || n_anchors == 0) return VN_OK;
  VN_REQUIRE(anchors_standup && anchors && gt_offsets && pos_equal_one && neg_equal_one && targets && workspace,
             "rpn_targets: null pointer");
  VN_REQUIRE(n_boxes_total == 0 || (gt_standup && gt_boxes), "rpn_targets: null box arrays");
  VN_REQUIRE(n_frames <= 65535, "rpn_targets: more than 65535 frames");
  VN_REQUIRE(((reinterpret_cast<uintptr_t>(anchors_standup) | reinterpret_cast<uintptr_t>(gt_standup) |
               reinterpret_cast<uintptr_t>(workspace)) & 15) == 0, "rpn_targets: 16-byte alignment required");
  VN_REQUIRE(workspace_bytes >= vn_rpn_targets_workspace_bytes(n_frames, n_anchors, n_boxes_total),
             "rpn_targets: workspace too small");
  cudaStream_t st = as_stream(stream);
  const int64_t n_cells = (int64_t)n_frames * n_anchors;
  unsigned long long* best = reinterpret_cast<unsigned long long*>(workspace);
  int32_t* pos_box = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(workspace) +
                                                align_up((size_t)(n_boxes_total > 0 ? n_boxes_total : 1) * 8, 16));
  int32_t* claim = pos_box + n_cells;
  const int64_t want = (n_cells + 255) / 256, cap = (int64_t)sm_count() * 8;
  k_targets_init<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(best, n_boxes_total, claim, n_cells);
  VN_LAUNCH_CHECK("k_targets_init");
  const dim3 grid((unsigned)((n_anchors + 255) / 256), (unsigned)n_frames);
  k_targets_scan<<<grid, 256, 0, st>>>(reinterpret_cast<const F4*>(anchors_standup), n_anchors,
                                       reinterpret_cast<const F4*>(gt_standup), gt_offsets, pos_iou, neg_iou, pos_box,
                                       neg_equal_one, best);
  VN_LAUNCH_CHECK("k_targets_scan");
  if (n_boxes_total > 0) {
    k_targets_claim<<<(unsigned)((n_boxes_total + 255) / 256), 256, 0, st>>>(gt_offsets, n_frames, n_anchors, best, claim);
    VN_LAUNCH_CHECK("k_targets_claim");
  }
  k_targets_write<<<grid, 256, 0, st>>>(anchors, n_anchors, gt_boxes, gt_offsets, anchor_h, pos_box, claim,
                                        pos_equal_one, targets);
  VN_LAUNCH_CHECK("k_targets_write");
  return VN_OK;
}

size_t vn_rpn_targets_static_capacity_bytes(void) { return kStaticScratchBytes; }

// vn_rpn_targets with its scratch in the library's own static device array: no allocation of any kind on this path
// (no cudaMalloc / cudaFree, no allocating library routine) - only the four launches of vn_rpn_targets.  Every call on
// a device uses the same array: k_targets_init re-arms the parts that must start from a known value, the rest is
// written before it is read, so stale contents of the previous call are harmless; two calls that overlap on the GPU race.
int vn_rpn_targets_static(const float* anchors_standup, const double* anchors, int32_t n_anchors, const float* gt_standup,
                          const double* gt_boxes, const int32_t* gt_offsets, int32_t n_frames, int32_t n_boxes_total,
                          float pos_iou, float neg_iou, double anchor_h, double* pos_equal_one, double* neg_equal_one,
                          double* targets, vn_stream_t stream) {
  VN_REQUIRE(n_frames >= 0 && n_anchors >= 0 && n_boxes_total >= 0, "rpn_targets_static: negative size");
  if (n_frames == 0 || n_anchors == 0) return VN_OK;
  const size_t need = vn_rpn_targets_workspace_bytes(n_frames, n_anchors, n_boxes_total);
  if (need > kStaticScratchBytes) {
    set_error("rpn_targets_static: %d frames x %d anchors with %d boxes need %zu bytes of scratch, the static array holds %zu "
              "(VN_RPN_TARGETS_STATIC_BYTES) - use vn_rpn_targets with a caller workspace", n_frames, n_anchors,
              n_boxes_total, need, kStaticScratchBytes);
    return VN_ERR_WORKSPACE;
  }
  void* scratch = nullptr;
  const int rc = static_scratch_for_current_device(&scratch);
  if (rc != VN_OK) return rc;
  return vn_rpn_targets(anchors_standup, anchors, n_anchors, gt_standup, gt_boxes, gt_offsets, n_frames, n_boxes_total,
                        pos_iou, neg_iou, anchor_h, pos_equal_one, neg_equal_one, targets, scratch, kStaticScratchBytes,
                        stream);
}

}  // extern "C"
