// b2rf (sm_100a) — error state, device queries, version.
#include "b2rf_internal.h"
#include "../../include/b2rf.h"

#include <string>

namespace b2rf {

static thread_local std::string g_err;
void set_error(const std::string& m) { g_err = m; }
const char* last_error_cstr() { return g_err.c_str(); }
bool cuda_ok(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return true;
  g_err = std::string(what) + ": " + cudaGetErrorString(e);
  return false;
}
int device_sm_count() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, dev) != cudaSuccess) return 148;
    cached = p.multiProcessorCount; cached_dev = dev;
  }
  return cached;
}

}  // namespace b2rf

extern "C" {

int b2rf_version(void) { return 100; }
const char* b2rf_last_error(void) { return b2rf::last_error_cstr(); }

}  // extern "C"
