// Host-side frame packer: a list of per-frame (N_i, 4) float32 arrays -> ONE contiguous (pinned) point buffer + CSR
// frame offsets, copied by a small pool of worker threads.
//
// This is the head of the reference's input path done the other way round: the reference voxelizes every frame on
// the host (voxelnet/dataset.py:51-58) and ships 12.85 MB of features per frame (voxelnet/model.py:302-303); here the
// raw points of a batch (1.95 MB per frame) are gathered into page-locked memory so that ONE asynchronous copy moves
// them to the device.  At 16 frames per 1.5 ms step that is a 20 GB/s host copy - more than one thread sustains, so
// the copy is split into 256 KB pieces over the pool.  Pure host code: no CUDA call in this file.
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <thread>
#include <vector>

#include "vn_common.cuh"

namespace vn {
namespace {

struct Piece {
  const char* src;
  char* dst;
  size_t bytes;
};

struct Job {
  std::vector<Piece> pieces;
  std::atomic<size_t> next{0};  // next piece to hand out
  int active = 0;               // workers currently copying pieces of this job (guarded by the pool mutex)
  int64_t ticket = 0;
};

// Workers take the first job in the queue that still has pieces to hand out, so the tail of one batch overlaps the
// head of the next; jobs are retired from the front in submission order (wait(ticket) = everything up to it is done).
class PackPool {
 public:
  explicit PackPool(int n_threads) {
    for (int i = 0; i < n_threads; ++i) workers_.emplace_back([this] { run(); });
  }
  ~PackPool() {
    {
      std::lock_guard<std::mutex> g(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : workers_) t.join();
  }
  int size() const { return (int)workers_.size(); }

  int64_t submit(Job* job) {
    std::lock_guard<std::mutex> g(mu_);
    job->ticket = ++last_ticket_;
    queue_.push_back(job);
    cv_.notify_all();
    return job->ticket;
  }

  void wait(int64_t ticket) {
    std::unique_lock<std::mutex> g(mu_);
    done_cv_.wait(g, [&] { return finished_ >= ticket; });
  }

 private:
  Job* pick() {  // mutex held
    for (Job* j : queue_)
      if (j->next.load(std::memory_order_relaxed) < j->pieces.size()) return j;
    return nullptr;
  }

  void run() {
    std::unique_lock<std::mutex> g(mu_);
    for (;;) {
      Job* job = nullptr;
      cv_.wait(g, [&] { return stop_ || (job = pick()) != nullptr; });
      if (job == nullptr) return;  // stop requested and nothing left to copy
      ++job->active;
      g.unlock();
      const size_t n = job->pieces.size();
      for (;;) {
        const size_t i = job->next.fetch_add(1, std::memory_order_relaxed);
        if (i >= n) break;
        const Piece& p = job->pieces[i];
        if (p.bytes != 0) std::memcpy(p.dst, p.src, p.bytes);
      }
      g.lock();
      --job->active;
      bool retired = false;
      while (!queue_.empty() && queue_.front()->active == 0 &&
             queue_.front()->next.load(std::memory_order_relaxed) >= queue_.front()->pieces.size()) {
        Job* f = queue_.front();
        queue_.pop_front();
        finished_ = f->ticket;
        delete f;
        retired = true;
      }
      if (retired) done_cv_.notify_all();
    }
  }

  std::vector<std::thread> workers_;
  std::deque<Job*> queue_;
  std::mutex mu_;
  std::condition_variable cv_, done_cv_;
  int64_t last_ticket_ = 0, finished_ = 0;
  bool stop_ = false;
};

std::mutex g_pool_mu;
PackPool* g_pool = nullptr;
int g_pool_threads = 0;

PackPool* pool() {
  std::lock_guard<std::mutex> g(g_pool_mu);
  if (g_pool == nullptr) {
    int n = g_pool_threads;
    if (n <= 0) {
      const char* e = std::getenv("VN_HOST_THREADS");
      n = e != nullptr ? std::atoi(e) : 4;
    }
    if (n < 1) n = 1;
    if (n > 64) n = 64;
    g_pool = new PackPool(n);
  }
  return g_pool;
}

constexpr size_t kPieceBytes = 256 * 1024;

}  // namespace
}  // namespace vn

using namespace vn;

extern "C" {

int vn_host_pack_threads(int32_t n_threads) {
  std::lock_guard<std::mutex> g(g_pool_mu);
  if (n_threads > 0 && g_pool == nullptr) g_pool_threads = n_threads;
  return g_pool != nullptr ? g_pool->size() : (g_pool_threads > 0 ? g_pool_threads : 0);
}

int64_t vn_host_pack_submit(const float* const* frames_host, const int64_t* n_points_host, int32_t n_frames,
                            float* points_host, int64_t capacity_points, int32_t* frame_offsets_host) {
  if (frames_host == nullptr || n_points_host == nullptr || points_host == nullptr || frame_offsets_host == nullptr ||
      n_frames < 0) {
    set_error("host_pack: null pointer or n_frames < 0");
    return VN_ERR_BAD_ARG;
  }
  int64_t total = 0;
  for (int i = 0; i < n_frames; ++i) {
    if (n_points_host[i] < 0 || (n_points_host[i] > 0 && frames_host[i] == nullptr)) {
      set_error("host_pack: frame %d has %lld points / null data", i, (long long)n_points_host[i]);
      return VN_ERR_BAD_ARG;
    }
    total += n_points_host[i];
  }
  if (total > capacity_points || total > 0x7fffffffLL) {
    set_error("host_pack: %lld points exceed the buffer capacity %lld", (long long)total, (long long)capacity_points);
    return VN_ERR_BAD_ARG;
  }
  Job* job = new Job();
  int64_t off = 0;
  frame_offsets_host[0] = 0;
  for (int i = 0; i < n_frames; ++i) {
    const char* src = reinterpret_cast<const char*>(frames_host[i]);
    char* dst = reinterpret_cast<char*>(points_host + off * 4);
    size_t left = (size_t)n_points_host[i] * 16;
    while (left > 0) {
      const size_t b = left < kPieceBytes ? left : kPieceBytes;
      job->pieces.push_back(Piece{src, dst, b});
      src += b; dst += b; left -= b;
    }
    off += n_points_host[i];
    frame_offsets_host[i + 1] = (int32_t)off;
  }
  if (job->pieces.empty()) job->pieces.push_back(Piece{nullptr, nullptr, 0});  // an empty batch still yields a ticket
  return pool()->submit(job);
}

int vn_host_pack_wait(int64_t ticket) {
  if (ticket <= 0) {
    set_error("host_pack_wait: ticket %lld", (long long)ticket);
    return VN_ERR_BAD_ARG;
  }
  pool()->wait(ticket);
  return VN_OK;
}

}  // extern "C"
