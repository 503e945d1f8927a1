// Library plumbing: version, last-error text, launch counter.
#include <stdarg.h>

#include <atomic>

#include "obst_common.cuh"

namespace obst {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};
static std::atomic<int64_t> g_paths[OBST_PATH_COUNT];

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
void count_path(int which) {
  if (which >= 0 && which < OBST_PATH_COUNT) g_paths[which].fetch_add(1, std::memory_order_relaxed);
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
    return OBST_ECUDA;
  }
  return OBST_OK;
}

}  // namespace obst

extern "C" {

int obst_version(void) { return 100; }

const char* obst_last_error(void) { return obst::g_err; }

int64_t obst_launch_count(void) { return obst::g_launches.load(std::memory_order_relaxed); }

void obst_launch_count_reset(void) { obst::g_launches.store(0, std::memory_order_relaxed); }

int64_t obst_path_count(int32_t which) {
  if (which < 0 || which >= OBST_PATH_COUNT) return -1;
  return obst::g_paths[which].load(std::memory_order_relaxed);
}

}  // extern "C"
