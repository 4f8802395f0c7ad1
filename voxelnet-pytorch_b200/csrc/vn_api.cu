// Library-wide host plumbing: error strings, launch accounting, grid constants.
#include <atomic>
#include <cstdarg>
#include <cstring>

#include "vn_common.cuh"

namespace vn {

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  return static_cast<int>(e);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

}  // namespace vn

extern "C" {

int vn_abi_version(void) { return VN_ABI_VERSION; }

const char* vn_last_error(void) { return vn::g_err; }

int64_t vn_launch_count(void) { return vn::g_launches.load(std::memory_order_relaxed); }

// Grid constants restated from voxelnet/utils.py:24-33: every value is a float32 there, so the
// decimal literals are rounded to float here the same way np.float32(0.2) is.
int vn_grid_for_target(const char* target, vn_grid_t* g) {
  VN_REQUIRE(target != nullptr && g != nullptr, "vn_grid_for_target: null argument");
  const bool car = std::strcmp(target, "Car") == 0;
  g->D = 10;
  g->H = car ? 400 : 200;
  g->W = car ? 352 : 240;
  g->T = car ? 35 : 45;
  g->off_x = 0.0f;
  g->off_y = car ? 40.0f : 20.0f;
  g->off_z = 3.0f;
  g->vs_x = 0.2f;
  g->vs_y = 0.2f;
  g->vs_z = 0.4f;
  return VN_OK;
}

// One batch of per-frame page-locked arrays -> the device CSR slot, no host-side packing: frame i goes to
// points_dev[offset_i ..] by its own cudaMemcpyAsync (the copy engine gathers; the host only enqueues), the offsets are
// written into the caller's page-locked frame_offsets_host and copied last.  Replaces the reference's per-element
// `.to(device)` of host-voxelized features (voxelnet/model.py:302-303) for loaders that read .bin files into pinned
// memory (dataset.load_velodyne_bin(pinned=True)).  All source arrays and frame_offsets_host must stay untouched until
// the stream has passed this point.
int vn_upload_frames(const float* const* frames_host, const int64_t* n_points_host, int32_t n_frames, float* points_dev,
                     int64_t capacity_points, int32_t* frame_offsets_host, int32_t* frame_offsets_dev, vn_stream_t stream) {
  VN_REQUIRE(n_frames >= 1 && frames_host && n_points_host && points_dev && frame_offsets_host && frame_offsets_dev,
             "upload_frames: null pointer or no frame");
  int64_t total = 0;
  for (int i = 0; i < n_frames; ++i) {
    VN_REQUIRE(n_points_host[i] >= 0 && (n_points_host[i] == 0 || frames_host[i] != nullptr), "upload_frames: frame %d is invalid", i);
    total += n_points_host[i];
  }
  VN_REQUIRE(total <= capacity_points && total < ((int64_t)1 << 31), "upload_frames: %lld points exceed the slot capacity %lld",
             (long long)total, (long long)capacity_points);
  cudaStream_t st = vn::as_stream(stream);
  int64_t off = 0;
  for (int i = 0; i < n_frames; ++i) {
    frame_offsets_host[i] = (int32_t)off;
    if (n_points_host[i] > 0)
      VN_CUDA(cudaMemcpyAsync(points_dev + off * 4, frames_host[i], (size_t)n_points_host[i] * 16, cudaMemcpyHostToDevice, st));
    off += n_points_host[i];
  }
  frame_offsets_host[n_frames] = (int32_t)off;
  VN_CUDA(cudaMemcpyAsync(frame_offsets_dev, frame_offsets_host, (size_t)(n_frames + 1) * 4, cudaMemcpyHostToDevice, st));
  return VN_OK;
}

}  // extern "C"
