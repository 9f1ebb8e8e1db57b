// Library-wide runtime support: launch accounting and error reporting.
#include <atomic>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"

__device__ unsigned long long dph_rng_epoch = 0;

__global__ void rng_epoch_kernel(unsigned long long value, int advance) {
  if (threadIdx.x == 0 && blockIdx.x == 0) dph_rng_epoch = advance ? dph_rng_epoch + 1ull : value;
}

__global__ void rng_epoch_read_kernel(unsigned long long* out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = dph_rng_epoch;
}

static std::atomic<long long> g_launches{0};
static thread_local char g_err[512] = "";

bool dph_pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("DPH_PDL");
    return !(e != nullptr && e[0] == '0');
  }();
  return on;
}

void dph_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

namespace dph {
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
}  // namespace dph

int dph_check_last(const char* what) {
  // cudaPeekAtLastError does not synchronise; it only reports launch-configuration errors.
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    dph::set_error("%s: %s", what, cudaGetErrorString(e));
    cudaGetLastError();
    return DPH_ERR_CUDA;
  }
  return DPH_OK;
}

extern "C" {
int dph_rng_set(unsigned long long value, void* stream) {
  rng_epoch_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(value, 0);
  dph_count_launch();
  return dph_check_last("rng_set");
}
int dph_rng_advance(void* stream) {
  rng_epoch_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(0ull, 1);
  dph_count_launch();
  return dph_check_last("rng_advance");
}
int dph_rng_get(void* out_dev, void* stream) {
  if (out_dev == nullptr) {
    dph::set_error("rng_get: null output pointer");
    return DPH_ERR_ARG;
  }
  rng_epoch_read_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<unsigned long long*>(out_dev));
  dph_count_launch();
  return dph_check_last("rng_get");
}
long long dph_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
void dph_launch_count_reset(void) { g_launches.store(0, std::memory_order_relaxed); }
int dph_abi_version(void) { return 1; }
const char* dph_last_error(void) { return g_err; }
}
