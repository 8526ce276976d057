// Device runtime: the B200-native replacement of dark_cuda.c (device select, allocation, push/pull,
// streams, error reporting).  Differences from the reference, on purpose:
//   * errors are returned as status codes + yc_last_error() instead of print-and-exit
//     (check_error, dark_cuda.c:58-83); the trampolines at the seam restore print-and-exit.
//   * no hidden per-device global stream (get_cuda_stream, dark_cuda.c:128-152): every entry point
//     takes its stream, so calls are re-entrant across devices and capturable into CUDA graphs.
//   * no cuBLAS / cuDNN handles: nothing on the path uses a library.
#include <atomic>
#include <string.h>

#include "yc_common.cuh"

namespace yc {

static thread_local char g_err[1024] = "";
static std::atomic<long long> g_launches{0};
static int g_sm_count = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int sm_count() {
  if (g_sm_count == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      g_sm_count = n;
    else
      g_sm_count = 148;
  }
  return g_sm_count;
}

}  // namespace yc

using namespace yc;

extern "C" const char* yc_last_error(void) { return g_err; }

extern "C" int yc_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

extern "C" int yc_set_device(int device) {
  int n = yc_device_count();
  if (n <= 0) {
    set_error("yc_set_device: no CUDA device visible; this library has no CPU fallback");
    return YC_ERR_NO_DEVICE;
  }
  YC_REQUIRE(device >= 0 && device < n, "yc_set_device: device %d out of range [0,%d)", device, n);
  YC_CUDA(cudaSetDevice(device));
  int major = 0, minor = 0;
  YC_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
  YC_CUDA(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, device));
  if (major != 10) {
    set_error("yc_set_device: device %d is sm_%d%d; this library is built for sm_100a only", device, major, minor);
    return YC_ERR_NO_DEVICE;
  }
  g_sm_count = 0;
  sm_count();
  return YC_OK;
}

extern "C" int yc_get_device(void) {
  int d = -1;
  if (cudaGetDevice(&d) != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return d;
}

extern "C" int yc_sm_count(void) { return sm_count(); }

extern "C" void* yc_malloc(size_t bytes) {
  void* p = nullptr;
  if (bytes == 0) bytes = 16;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) {
    set_error("yc_malloc(%zu): %s", bytes, cudaGetErrorString(e));
    return nullptr;
  }
  e = cudaMemset(p, 0, bytes);
  if (e != cudaSuccess) {
    set_error("yc_malloc memset: %s", cudaGetErrorString(e));
    cudaFree(p);
    return nullptr;
  }
  return p;
}
extern "C" int yc_free(void* p) {
  if (p) YC_CUDA(cudaFree(p));
  return YC_OK;
}
extern "C" void* yc_malloc_host(size_t bytes) {
  void* p = nullptr;
  cudaError_t e = cudaHostAlloc(&p, bytes ? bytes : 16, cudaHostAllocDefault);
  if (e != cudaSuccess) {
    set_error("yc_malloc_host(%zu): %s", bytes, cudaGetErrorString(e));
    return nullptr;
  }
  return p;
}
extern "C" int yc_free_host(void* p) {
  if (p) YC_CUDA(cudaFreeHost(p));
  return YC_OK;
}
extern "C" int yc_push(void* dst, const void* src, size_t bytes, void* stream) {
  YC_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return YC_OK;
}
extern "C" int yc_pull(void* dst, const void* src, size_t bytes, void* stream) {
  YC_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return YC_OK;
}
extern "C" int yc_memset(void* dst, int value, size_t bytes, void* stream) {
  YC_CUDA(cudaMemsetAsync(dst, value, bytes, (cudaStream_t)stream));
  return YC_OK;
}
extern "C" int yc_stream_sync(void* stream) {
  YC_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return YC_OK;
}
extern "C" int yc_stream_create(void** stream) {
  cudaStream_t s;
  YC_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *stream = (void*)s;
  return YC_OK;
}
extern "C" int yc_stream_destroy(void* stream) {
  YC_CUDA(cudaStreamDestroy((cudaStream_t)stream));
  return YC_OK;
}
extern "C" long long yc_launch_count(int reset) {
  long long v = g_launches.load();
  if (reset) g_launches.store(0);
  return v;
}
