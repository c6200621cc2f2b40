// Library-level entry points: version, thread-local error string, device info.
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "common.cuh"

static thread_local char g_err[512] = "";

void b2f_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int b2f_sm_count() {
  static int cached[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

cudaError_t b2f_smem_attr_impl(const void* kernel, size_t bytes) {
  struct Entry {
    const void* fn;
    int dev;
    size_t bytes;
  };
  static Entry table[256];
  static int used = 0;
  static std::mutex mu;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lock(mu);
  Entry* hit = nullptr;
  for (int i = 0; i < used; ++i)
    if (table[i].fn == kernel && table[i].dev == dev) hit = &table[i];
  if (hit && hit->bytes >= bytes) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) return e;
  if (hit) hit->bytes = bytes;
  else if (used < 256) table[used++] = Entry{kernel, dev, bytes};
  return cudaSuccess;
}

extern "C" int b2f_version(void) { return 200; }

extern "C" const char* b2f_last_error(void) { return g_err; }
