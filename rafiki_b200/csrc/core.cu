// Error reporting, version and device checks of librafiki_b200.
#include <stdarg.h>
#include <string.h>

#include <atomic>

#include "rfk_common.cuh"

namespace rfk {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// Launch accounting: one counter per kernel name (the string literals passed to RFK_LAUNCH_CHECK).
// A slot is claimed with a compare-exchange on its name pointer; counting is one relaxed add.
namespace {
constexpr int kMaxKernelNames = 64;
std::atomic<const char*> g_kernel_name[kMaxKernelNames];
std::atomic<unsigned long long> g_kernel_launches[kMaxKernelNames];
}  // namespace

void count_launch(const char* name) {
  for (int i = 0; i < kMaxKernelNames; ++i) {
    const char* cur = g_kernel_name[i].load(std::memory_order_acquire);
    if (cur == nullptr) {
      const char* expected = nullptr;
      if (g_kernel_name[i].compare_exchange_strong(expected, name, std::memory_order_acq_rel)) cur = name;
      else cur = expected;
    }
    if (cur == name || strcmp(cur, name) == 0) {
      g_kernel_launches[i].fetch_add(1, std::memory_order_relaxed);
      return;
    }
  }
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

}  // namespace rfk

extern "C" int rfk_version(void) { return RFK_VERSION; }

extern "C" const char* rfk_last_error(void) { return rfk::g_err; }

extern "C" unsigned long long rfk_launch_count(const char* name_part) {
  unsigned long long total = 0;
  for (int i = 0; i < rfk::kMaxKernelNames; ++i) {
    const char* nm = rfk::g_kernel_name[i].load(std::memory_order_acquire);
    if (nm == nullptr) break;
    if (name_part == nullptr || name_part[0] == 0 || strstr(nm, name_part) != nullptr)
      total += rfk::g_kernel_launches[i].load(std::memory_order_relaxed);
  }
  return total;
}

extern "C" int rfk_device_check(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    rfk::set_error("no CUDA device: %s (this library has no CPU fallback)",
                   e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    return RFK_ERR_CUDA;
  }
  RFK_REQUIRE(device >= 0 && device < count, "device %d out of range (have %d)", device, count);
  int major = 0, minor = 0;
  RFK_CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  RFK_CUDA_TRY(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  if (major != 10) {
    rfk::set_error("device %d is sm_%d%d; librafiki_b200 is built for sm_100a only", device, major,
                   minor);
    return RFK_ERR_UNSUPPORTED;
  }
  return RFK_OK;
}
