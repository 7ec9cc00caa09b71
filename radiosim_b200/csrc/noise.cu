// add_noise (reference src/radiosim/utils/utils.py:259-315; hooks simulations/simulations.py:101-108,
// 121-125): per bundle a white field plus two Gaussian-filtered fields, each scaled by its
// largest magnitude over the WHOLE bundle, summed, scaled per image and added to the image.
//
// The fields are the reference's own draws: the host replays the legacy np.random stream
// (3 n N^2 normals, rsb_legacy_normal_fill) and ships them as float32.  The reference filters with
// scipy.signal.fftconvolve(mode="same") against astropy's Gaussian2DKernel(kernel / 2) -- an outer
// product, so here: two zero-padded 1-D passes, taps g(k) = exp(-k^2 / 2 sigma^2) / sum,
// |k| <= h, 2h + 1 = 8 sigma rounded up to an odd integer.
//
//   hblur_kernel   : row pass of both filtered fields (x max_noise of the image), row staged in
//                    shared memory with its zero halo
//   vblur_kernel   : column pass from a shared-memory tile; largest magnitude of each field ->
//                    atomicMax
//   absmax_kernel  : largest magnitude of the white field
//   combine_kernel : 0.2 W/max|W| + 0.3 A/max|A| + 0.5 B/max|B|; per-image maxima of the sum and of
//                    the image
//   final_kernel   : image + sum / max(sum) * max(image) * level        (float64)
#include "common.cuh"

namespace rsb {
namespace noise {

constexpr int kMaxTaps = 1025;

struct Taps {  // one filtered field
  float sigma;
  int h;
};

__host__ __device__ inline int half_width(double sigma) {
  int size = (int)ceil(8.0 * sigma);
  if ((size & 1) == 0) size += 1;
  return (size - 1) / 2;
}

// normalised taps into shared memory (every thread of the block calls this)
__device__ inline void build_taps(float* g, float sigma, int h) {
  __shared__ float s_sum;
  if (threadIdx.x == 0 && threadIdx.y == 0) {
    double sum = 0.0;
    for (int k = -h; k <= h; ++k) sum += exp(-(double)k * k / (2.0 * (double)sigma * sigma));
    s_sum = (float)(1.0 / sum);
  }
  __syncthreads();
  const int t = threadIdx.y * blockDim.x + threadIdx.x, nt = blockDim.x * blockDim.y;
  for (int k = t; k <= 2 * h; k += nt) {
    const double d = (double)(k - h);
    g[k] = (float)(exp(-d * d / (2.0 * (double)sigma * sigma))) * s_sum;
  }
  __syncthreads();
}

// grid (N rows, n images, 2 fields), block 256.  smem: taps [2h+1] + row [N + 2h]
__global__ void __launch_bounds__(256) hblur_kernel(const float* __restrict__ fa, const float* __restrict__ fb,
                                                    const float* __restrict__ amp_a, const float* __restrict__ amp_b,
                                                    int N, Taps ta, Taps tb, float* __restrict__ tmp, size_t field_px) {
  extern __shared__ float sm[];
  const int f = blockIdx.z, img = blockIdx.y, r = blockIdx.x;
  const Taps t = f ? tb : ta;
  float* g = sm;
  float* row = sm + 2 * t.h + 1;
  build_taps(g, t.sigma, t.h);
  const float* src = (f ? fb : fa) + ((size_t)img * N + r) * N;
  const float amp = (f ? amp_b : amp_a)[img];
  for (int c = threadIdx.x; c < N + 2 * t.h; c += blockDim.x) {
    const int cc = c - t.h;
    row[c] = (cc >= 0 && cc < N) ? src[cc] * amp : 0.f;
  }
  __syncthreads();
  float* dst = tmp + (size_t)f * field_px + ((size_t)img * N + r) * N;
  for (int c = threadIdx.x; c < N; c += blockDim.x) {
    float s = 0.f;
    for (int k = 0; k <= 2 * t.h; ++k) s = fmaf(g[k], row[c + k], s);
    dst[c] = s;
  }
}

// grid (ceil(N/32) column blocks, n images, 2 fields), block (32, 8).
// smem: taps [2h+1] + tile [N + 2h][32]
__global__ void __launch_bounds__(256) vblur_kernel(const float* __restrict__ tmp, int N, Taps ta, Taps tb,
                                                    float* __restrict__ out, size_t field_px,
                                                    unsigned* __restrict__ absmax) {
  extern __shared__ float sm[];
  const int f = blockIdx.z, img = blockIdx.y, c = blockIdx.x * 32 + threadIdx.x;
  const Taps t = f ? tb : ta;
  float* g = sm;
  float* tile = sm + ((2 * t.h + 1 + 31) & ~31);
  build_taps(g, t.sigma, t.h);
  const float* src = tmp + (size_t)f * field_px + (size_t)img * N * N;
  for (int rr = threadIdx.y; rr < N + 2 * t.h; rr += blockDim.y) {
    const int r = rr - t.h;
    tile[rr * 32 + threadIdx.x] = (r >= 0 && r < N && c < N) ? src[(size_t)r * N + c] : 0.f;
  }
  __syncthreads();
  float* dst = out + (size_t)f * field_px + (size_t)img * N * N;
  float m = 0.f;
  if (c < N) {
    for (int r = threadIdx.y; r < N; r += blockDim.y) {
      float s = 0.f;
      for (int k = 0; k <= 2 * t.h; ++k) s = fmaf(g[k], tile[(r + k) * 32 + threadIdx.x], s);
      dst[(size_t)r * N + c] = s;
      m = fmaxf(m, fabsf(s));
    }
  }
  m = warp_max(m);
  if (threadIdx.x == 0) atomicMax(absmax + 1 + f, __float_as_uint(m));  // (|.| >= 0: float order = uint order)
}

__global__ void __launch_bounds__(256) absmax_kernel(const float* __restrict__ w, size_t count,
                                                     unsigned* __restrict__ absmax) {
  float m = 0.f;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < count; e += (size_t)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf(w[e]));
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) atomicMax(absmax, __float_as_uint(m));
}

// order-preserving map of a float / double to an unsigned integer (for atomicMax on signed values)
__device__ __forceinline__ unsigned f32_key(float v) {
  const unsigned b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float f32_unkey(unsigned k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
__device__ __forceinline__ unsigned long long f64_key(double v) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double f64_unkey(unsigned long long k) {
  const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
  return __longlong_as_double((long long)b);
}

// grid (pixel blocks, n images), block 256
__global__ void __launch_bounds__(256) combine_kernel(const float* __restrict__ white, const float* __restrict__ blurred,
                                                      size_t field_px, const double* __restrict__ image, int N,
                                                      const unsigned* __restrict__ absmax, float* __restrict__ sum,
                                                      unsigned* __restrict__ sum_max,
                                                      unsigned long long* __restrict__ img_max) {
  const int img = blockIdx.y;
  const size_t npx = (size_t)N * N, base = (size_t)img * npx;
  const float s0 = 0.2f / __uint_as_float(absmax[0]), s1 = 0.3f / __uint_as_float(absmax[1]),
              s2 = 0.5f / __uint_as_float(absmax[2]);
  float m = -INFINITY;
  double im = -INFINITY;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < npx; e += (size_t)gridDim.x * blockDim.x) {
    // the reference's order of additions: white, then the N/32 field, then the N/8 field
    const float v = (white[base + e] * s0 + blurred[base + e] * s1) + blurred[field_px + base + e] * s2;
    sum[base + e] = v;
    m = fmaxf(m, v);
    im = fmax(im, image[base + e]);
  }
  m = warp_max(m);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) im = fmax(im, __shfl_xor_sync(0xffffffffu, im, o));
  if ((threadIdx.x & 31) == 0) {
    atomicMax(sum_max + img, f32_key(m));
    atomicMax(img_max + img, f64_key(im));
  }
}

__global__ void __launch_bounds__(256) final_kernel(const double* __restrict__ image, const float* __restrict__ sum,
                                                    const unsigned* __restrict__ sum_max,
                                                    const unsigned long long* __restrict__ img_max,
                                                    const double* __restrict__ level, int N, double* __restrict__ out) {
  const int img = blockIdx.y;
  const size_t npx = (size_t)N * N, base = (size_t)img * npx;
  const double scale = f64_unkey(img_max[img]) * level[img] / (double)f32_unkey(sum_max[img]);
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < npx; e += (size_t)gridDim.x * blockDim.x)
    out[base + e] = (double)sum[base + e] * scale + image[base + e];
}

}  // namespace noise
}  // namespace rsb

using namespace rsb;
using namespace rsb::noise;

static size_t align256n(size_t b) { return (b + 255) & ~(size_t)255; }

extern "C" {

// workspace: [maxima: |white|, |A|, |B| (3 x u32) | per-image max of the sum (n x u32) | per-image max
// of the image (n x u64)] then four float32 fields (row pass x 2, filtered x 2)
static size_t noise_ctl_bytes(int n, size_t* off_summax, size_t* off_imgmax) {
  *off_summax = 64;
  *off_imgmax = 64 + (((size_t)n * 4 + 7) & ~(size_t)7);
  return align256n(*off_imgmax + (size_t)n * 8);
}

int rsb_noise_workspace_bytes(int n, int size, size_t* bytes) {
  RSB_ARG(bytes && n >= 0 && size >= 2, "rsb_noise_workspace_bytes: bad arguments");
  const int m = n > 0 ? n : 1;
  size_t a, b;
  *bytes = noise_ctl_bytes(m, &a, &b) + 4 * align256n((size_t)m * size * size * sizeof(float));
  return 0;
}

int rsb_add_noise(const double* image, const float* white, const float* field_a, const float* field_b,
                  const float* amp_a, const float* amp_b, const double* level, int n, int size, double* out,
                  void* workspace, size_t workspace_bytes, void* stream) {
  RSB_ARG(n >= 0 && size >= 2, "rsb_add_noise: bad n/size (%d, %d)", n, size);
  if (n == 0) return 0;
  RSB_ARG(n <= 65535, "rsb_add_noise: at most 65535 images per call");
  RSB_ARG(image && white && field_a && field_b && amp_a && amp_b && level && out && workspace,
          "rsb_add_noise: NULL buffer");
  size_t need = 0;
  if (int e = rsb_noise_workspace_bytes(n, size, &need)) return e;
  RSB_ARG(workspace_bytes >= need, "rsb_add_noise: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
  // kernels = [1, N/32, N/8], Gaussian2DKernel(kernel / 2)  (utils.py:281, :301)
  Taps ta, tb;
  ta.sigma = (float)((double)size / 32.0 / 2.0); ta.h = half_width((double)size / 32.0 / 2.0);
  tb.sigma = (float)((double)size / 8.0 / 2.0);  tb.h = half_width((double)size / 8.0 / 2.0);
  const size_t smem_h = (size_t)(2 * tb.h + 1 + size + 2 * tb.h) * sizeof(float);
  const size_t smem_v = (size_t)(((2 * tb.h + 1 + 31) & ~31) + (size + 2 * tb.h) * 32) * sizeof(float);
  RSB_ARG(2 * tb.h + 1 <= kMaxTaps && smem_v <= 227 * 1024,
          "rsb_add_noise: image size %d needs %zu bytes of shared memory (limit 227 KiB: size <= 1024)", size, smem_v);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t px = (size_t)n * size * size, fbytes = align256n(px * sizeof(float));
  char* w = (char*)workspace;
  size_t off_summax, off_imgmax;
  const size_t off = noise_ctl_bytes(n, &off_summax, &off_imgmax);
  unsigned* absmax = (unsigned*)w;                                        // [3]
  unsigned* sum_max = (unsigned*)(w + off_summax);                        // [n]
  unsigned long long* img_max = (unsigned long long*)(w + off_imgmax);    // [n]
  float* tmp = (float*)(w + off);            // 2 fields: row pass
  float* blurred = tmp + 2 * (fbytes / 4);   // 2 fields: both passes
  const size_t field_px = fbytes / 4;
  RSB_CUDA(cudaMemsetAsync(w, 0, off, st));
  RSB_CUDA(cudaFuncSetAttribute(hblur_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_h));
  RSB_CUDA(cudaFuncSetAttribute(vblur_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_v));
  hblur_kernel<<<dim3(size, n, 2), 256, smem_h, st>>>(field_a, field_b, amp_a, amp_b, size, ta, tb, tmp, field_px);
  RSB_LAUNCHED();
  vblur_kernel<<<dim3((size + 31) / 32, n, 2), dim3(32, 8), smem_v, st>>>(tmp, size, ta, tb, blurred, field_px, absmax);
  RSB_LAUNCHED();
  absmax_kernel<<<sm_count() * 4, 256, 0, st>>>(white, px, absmax);
  RSB_LAUNCHED();
  int bx = (int)(((size_t)size * size + 255) / 256);
  if (bx > 64) bx = 64;
  float* sum = tmp;  // the row-pass scratch is free again
  combine_kernel<<<dim3(bx, n), 256, 0, st>>>(white, blurred, field_px, image, size, absmax, sum, sum_max, img_max);
  RSB_LAUNCHED();
  final_kernel<<<dim3(bx, n), 256, 0, st>>>(image, sum, sum_max, img_max, level, size, out);
  RSB_LAUNCHED();
  return 0;
}

}  // extern "C"
