// Status plumbing and device queries shared by every entry point of libaiko_b200.so.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "common.cuh"

namespace aiko {

static thread_local char g_last_error[512] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
}

int cuda_status(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_last_error("%s: %s (%s)", what, cudaGetErrorString(e), cudaGetErrorName(e));
  return AIKO_B200_ERR_CUDA;
}

// per-device attribute cache (a process may drive several GPUs from worker threads)
static int g_sm[64];
static int g_smem[64];

static int cur_dev() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= 64) d = 0;
  return d;
}

int sm_count() {
  const int d = cur_dev();
  if (g_sm[d] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, d) != cudaSuccess || v <= 0) v = 148;
    g_sm[d] = v;
  }
  return g_sm[d];
}

int device_max_smem_optin() {
  const int d = cur_dev();
  if (g_smem[d] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, d) != cudaSuccess || v <= 0)
      v = 227 * 1024;
    g_smem[d] = v;
  }
  return g_smem[d];
}

}  // namespace aiko

extern "C" int aiko_b200_abi_version(void) { return AIKO_B200_ABI_VERSION; }

extern "C" const char* aiko_b200_last_error(void) { return aiko::g_last_error; }

extern "C" int aiko_b200_device_info(char* name, size_t name_len, int* cc, int* sms) {
  int dev = 0;
  AIKO_CUDA(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  AIKO_CUDA(cudaGetDeviceProperties(&prop, dev));
  if (name && name_len) {
    strncpy(name, prop.name, name_len - 1);
    name[name_len - 1] = 0;
  }
  if (cc) *cc = prop.major * 10 + prop.minor;
  if (sms) *sms = prop.multiProcessorCount;
  return AIKO_B200_OK;
}
