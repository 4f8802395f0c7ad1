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

}  // extern "C"
