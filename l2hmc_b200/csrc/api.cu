// C-ABI glue shared by all entry points: error text, version, device limits.
#include <stdarg.h>
#include <stdlib.h>

#include <atomic>

#include "common.cuh"

namespace l2hmc {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

// Every kernel launch of the library is followed by one check_launch (with the number of launches it covers): the
// process-wide tally is what bench.py reports as `gpu_launches` (counted where the launches happen, not estimated).
static std::atomic<unsigned long long> g_launches{0};

int check_launch(const char* what, int launches) {
  g_launches.fetch_add(static_cast<unsigned long long>(launches), std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return L2HMC_ERR_CUDA;
  }
  return L2HMC_OK;
}

// L2HMC_PDL=0: launch every kernel without the programmatic-dependent-launch attribute (diagnostics)
bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("L2HMC_PDL");
    return !(e && e[0] == '0' && e[1] == 0);
  }();
  return on;
}

}  // namespace l2hmc

extern "C" int l2hmc_abi_version(void) { return L2HMC_ABI_VERSION; }

extern "C" const char* l2hmc_last_error(void) { return l2hmc::g_error; }

extern "C" unsigned long long l2hmc_launch_count(void) { return l2hmc::g_launches.load(std::memory_order_relaxed); }

extern "C" int l2hmc_device_info(int* sm_count, int* max_smem_per_block) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    l2hmc::set_error("cudaGetDevice: %s", cudaGetErrorString(e));
    return L2HMC_ERR_CUDA;
  }
  if (sm_count) cudaDeviceGetAttribute(sm_count, cudaDevAttrMultiProcessorCount, dev);
  if (max_smem_per_block) cudaDeviceGetAttribute(max_smem_per_block, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  return L2HMC_OK;
}
