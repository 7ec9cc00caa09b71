// Device micro-benchmarks that give the roofline denominators bench.py reports next
// to the measured HBM copy figure: the MUFU.EX2 issue rate (the governing pipe of the
// exp-bound splat kernels).
#include "common.cuh"

namespace rsb {

__global__ void __launch_bounds__(256) sfu_kernel(float* sink, int iters, float seed) {
  // 8 independent chains per thread keep the SFU queue full; the FMUL that feeds each
  // EX2 runs on the FMA pipe, which has 8x the SFU issue rate.
  float x[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) x[j] = seed + 1e-3f * (float)(threadIdx.x + 32 * j);
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) x[j] = ex2(x[j] * -0.5f);
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += x[j];
  if (s == 12345.678f) sink[0] = s;  // never true: keeps the chains alive
}

}  // namespace rsb

extern "C" int rsb_bench_sfu(int iters, double* ex2_per_second) {
  using namespace rsb;
  RSB_ARG(ex2_per_second != nullptr && iters > 0, "rsb_bench_sfu: bad arguments");
  float* sink = nullptr;
  RSB_CUDA(cudaMalloc(&sink, 16));
  cudaEvent_t e0, e1;
  RSB_CUDA(cudaEventCreate(&e0));
  RSB_CUDA(cudaEventCreate(&e1));
  const int grid = sm_count() * 8, block = 256;
  sfu_kernel<<<grid, block>>>(sink, iters / 8 + 1, 0.25f);  // warm-up
  RSB_LAUNCHED();
  double best = 0.0;
  for (int rep = 0; rep < 5; ++rep) {
    RSB_CUDA(cudaEventRecord(e0));
    sfu_kernel<<<grid, block>>>(sink, iters, 0.25f);
    RSB_LAUNCHED();
    RSB_CUDA(cudaEventRecord(e1));
    RSB_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    RSB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    const double rate = (double)grid * block * 8.0 * (double)iters / ((double)ms * 1e-3);
    if (rate > best) best = rate;
  }
  RSB_CUDA(cudaEventDestroy(e0));
  RSB_CUDA(cudaEventDestroy(e1));
  RSB_CUDA(cudaFree(sink));
  *ex2_per_second = best;
  return 0;
}
