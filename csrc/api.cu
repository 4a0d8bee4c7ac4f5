// Library-level entry points: error string, version, device query, bucket planner.
#include "common.cuh"
#include "../include/mlt_b200.h"
#include <cstring>
#include <vector>

namespace mlt {
static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error() { return g_err; }
}  // namespace mlt

extern "C" {

const char* mlt_last_error(void) { return mlt::last_error(); }

int mlt_version(void) { return MLT_B200_VERSION; }

int mlt_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, long long* total_mem) {
  cudaDeviceProp p;
  MLT_CHECK_CUDA(cudaGetDeviceProperties(&p, device));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (total_mem) *total_mem = (long long)p.totalGlobalMem;
  return 0;
}

// Greedy contiguous bucket plan over parameters that are already in gradient-ready order.
// out_bounds receives nb+1 indices: bucket i covers params [out_bounds[i], out_bounds[i+1]).
// A bucket closes as soon as its byte count reaches the cap (first bucket has its own, smaller cap,
// like DDP's 1 MiB first bucket, so communication starts early in backward).
int mlt_plan_buckets(const long long* sizes_bytes, int n, long long cap_bytes, long long first_cap_bytes,
                     int* out_bounds) {
  int nb = 0;
  long long acc = 0, cap = first_cap_bytes > 0 ? first_cap_bytes : cap_bytes;
  out_bounds[0] = 0;
  for (int i = 0; i < n; ++i) {
    acc += sizes_bytes[i];
    if (acc >= cap) {
      out_bounds[++nb] = i + 1;
      acc = 0;
      cap = cap_bytes;
    }
  }
  if (out_bounds[nb] < n) out_bounds[++nb] = n;
  return nb;
}

}  // extern "C"
