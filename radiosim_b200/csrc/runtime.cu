// Runtime plumbing of the C ABI: errors, memory, streams, events.
#include <cstdarg>
#include <cstring>

#include "common.cuh"

namespace rsb {

static thread_local char t_err[512] = "";
std::atomic<uint64_t> g_launches{0};

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(t_err, sizeof(t_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
  const char* base = strrchr(file, '/');
  set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e),
            base ? base + 1 : file, line, what);
  return -(int)e;
}

int sm_count() {
  static int cached[64];
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

}  // namespace rsb

extern "C" {

int rsb_abi_version(void) { return RSB_ABI_VERSION; }
const char* rsb_last_error(void) { return rsb::t_err; }

int rsb_device_count(int* count) {
  RSB_ARG(count != nullptr, "rsb_device_count: count is NULL");
  RSB_CUDA(cudaGetDeviceCount(count));
  return 0;
}
int rsb_set_device(int device) {
  RSB_CUDA(cudaSetDevice(device));
  return 0;
}
int rsb_device_info(int device, int* sms, int* major, int* minor, size_t* total_mem) {
  cudaDeviceProp p;
  RSB_CUDA(cudaGetDeviceProperties(&p, device));
  if (sms) *sms = p.multiProcessorCount;
  if (major) *major = p.major;
  if (minor) *minor = p.minor;
  if (total_mem) *total_mem = p.totalGlobalMem;
  return 0;
}
int rsb_malloc(void** p, size_t bytes) {
  RSB_ARG(p != nullptr, "rsb_malloc: NULL out pointer");
  RSB_CUDA(cudaMalloc(p, bytes ? bytes : 1));
  return 0;
}
int rsb_free(void* p) {
  RSB_CUDA(cudaFree(p));
  return 0;
}
int rsb_malloc_host(void** p, size_t bytes) {
  RSB_ARG(p != nullptr, "rsb_malloc_host: NULL out pointer");
  RSB_CUDA(cudaMallocHost(p, bytes ? bytes : 1));
  return 0;
}
int rsb_free_host(void* p) {
  RSB_CUDA(cudaFreeHost(p));
  return 0;
}
int rsb_memset(void* p, int value, size_t bytes, void* stream) {
  RSB_CUDA(cudaMemsetAsync(p, value, bytes, (cudaStream_t)stream));
  return 0;
}
int rsb_memcpy_h2d(void* dst, const void* src, size_t bytes, void* stream) {
  RSB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return 0;
}
int rsb_memcpy_d2h(void* dst, const void* src, size_t bytes, void* stream) {
  RSB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return 0;
}
int rsb_memcpy2d_d2h(void* dst, size_t dst_pitch, const void* src, size_t src_pitch, size_t width_bytes,
                     size_t rows, void* stream) {
  RSB_CUDA(cudaMemcpy2DAsync(dst, dst_pitch, src, src_pitch, width_bytes, rows, cudaMemcpyDeviceToHost,
                             (cudaStream_t)stream));
  return 0;
}
int rsb_get_device(int* device) {
  RSB_ARG(device != nullptr, "rsb_get_device: NULL out pointer");
  RSB_CUDA(cudaGetDevice(device));
  return 0;
}
// Test aid: every SM's shared memory is left holding `pattern` (kernels must not depend on what
// the previous kernel left there; the parity tests run the renderers behind this).
__global__ void scrub_shared_kernel(unsigned pattern, int words) {
  extern __shared__ unsigned scrub_words[];
  for (int i = threadIdx.x; i < words; i += blockDim.x) scrub_words[i] = pattern;
  __syncthreads();
  if (scrub_words[(threadIdx.x * 97) % words] != pattern) __trap();  // (keeps the stores alive)
}
int rsb_debug_scrub_shared(unsigned pattern, void* stream) {
  int dev = 0, optin = 0, sms = 0;
  RSB_CUDA(cudaGetDevice(&dev));
  RSB_CUDA(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  RSB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  RSB_CUDA(cudaFuncSetAttribute(scrub_shared_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, optin));
  // one CTA per SM at a time (the whole opt-in size each); several waves so that none is missed
  scrub_shared_kernel<<<4 * sms, 256, optin, (cudaStream_t)stream>>>(pattern, optin / 4);
  RSB_CUDA(cudaGetLastError());
  return 0;
}
int rsb_stream_create(void** stream) {
  RSB_ARG(stream != nullptr, "rsb_stream_create: NULL out pointer");
  cudaStream_t s;
  RSB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *stream = (void*)s;
  return 0;
}
int rsb_stream_destroy(void* stream) {
  RSB_CUDA(cudaStreamDestroy((cudaStream_t)stream));
  return 0;
}
int rsb_stream_sync(void* stream) {
  RSB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
  return 0;
}
int rsb_device_sync(void) {
  RSB_CUDA(cudaDeviceSynchronize());
  return 0;
}
int rsb_event_create(void** event) {
  RSB_ARG(event != nullptr, "rsb_event_create: NULL out pointer");
  cudaEvent_t e;
  RSB_CUDA(cudaEventCreate(&e));
  *event = (void*)e;
  return 0;
}
int rsb_event_destroy(void* event) {
  RSB_CUDA(cudaEventDestroy((cudaEvent_t)event));
  return 0;
}
int rsb_event_record(void* event, void* stream) {
  RSB_CUDA(cudaEventRecord((cudaEvent_t)event, (cudaStream_t)stream));
  return 0;
}
int rsb_stream_wait_event(void* stream, void* event) {
  RSB_CUDA(cudaStreamWaitEvent((cudaStream_t)stream, (cudaEvent_t)event, 0));
  return 0;
}
int rsb_event_elapsed_ms(void* start, void* stop, float* ms) {
  RSB_ARG(ms != nullptr, "rsb_event_elapsed_ms: NULL out pointer");
  RSB_CUDA(cudaEventSynchronize((cudaEvent_t)stop));
  RSB_CUDA(cudaEventElapsedTime(ms, (cudaEvent_t)start, (cudaEvent_t)stop));
  return 0;
}
int rsb_launch_count(uint64_t* count) {
  RSB_ARG(count != nullptr, "rsb_launch_count: NULL out pointer");
  *count = rsb::g_launches.load();
  return 0;
}

}  // extern "C"
