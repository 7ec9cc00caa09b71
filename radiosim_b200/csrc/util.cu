// Stand-alone device versions of the image utilities the reference exports
// (src/radiosim/utils/utils.py): the threshold margin of zoom_on_source (:111-119) and
// cv2.resize(..., INTER_LINEAR) with optional zero padding (zoom_on_source :124-135,
// zoom_out :161-166), float64 in / float64 out like the reference.  They serve the
// drop-in utils.zoom_on_source / utils.zoom_out; the jet kernel has its own fused form.
#include "common.cuh"

namespace rsb {
namespace util {

__global__ void margin_kernel(const double* __restrict__ img, int N, double thr, int* margin) {
  int best = N;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < N * N; e += gridDim.x * blockDim.x) {
    if (img[e] > thr) {
      const int r = e / N, c = e - r * N;
      best = min(best, min(min(r, N - 1 - r), min(c, N - 1 - c)));
    }
  }
  best = warp_min(best);
  if ((threadIdx.x & 31) == 0 && best < N) atomicMin(margin, best);
}

// dst[D x D] = resize(pad(src[S x S] (row pitch `pitch`), pad)), cv2 INTER_LINEAR semantics:
// f = (d + 0.5) * P / D - 0.5, s = floor(f), w = f - s, clamped to [0, P-1] with w = 0.
__device__ __forceinline__ void taps(int d, int P, int D, int* s0, int* s1, double* w) {
  const double f = (d + 0.5) * ((double)P / (double)D) - 0.5;
  int s = (int)floor(f);
  double ww = f - (double)s;
  if (s < 0) { s = 0; ww = 0.0; }
  if (s >= P - 1) { s = P - 1; ww = 0.0; }
  *s0 = s;
  *s1 = min(s + 1, P - 1);
  *w = ww;
}

__global__ void resize_kernel(const double* __restrict__ src, int S, int pitch, int pad,
                              double* __restrict__ dst, int D, size_t src_plane, size_t dst_plane) {
  const int P = S + 2 * pad;
  const double* sp = src + (size_t)blockIdx.y * src_plane;
  double* dp = dst + (size_t)blockIdx.y * dst_plane;
  auto at = [&](int r, int c) -> double {
    r -= pad; c -= pad;
    return (r >= 0 && r < S && c >= 0 && c < S) ? sp[(size_t)r * pitch + c] : 0.0;
  };
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < D * D; e += gridDim.x * blockDim.x) {
    const int r = e / D, c = e - r * D;
    int r0, r1, c0, c1;
    double wr, wc;
    taps(r, P, D, &r0, &r1, &wr);
    taps(c, P, D, &c0, &c1, &wc);
    // horizontal pass first, then vertical (cv2)
    const double top = at(r0, c0) * (1.0 - wc) + at(r0, c1) * wc;
    const double bot = at(r1, c0) * (1.0 - wc) + at(r1, c1) * wc;
    dp[e] = top * (1.0 - wr) + bot * wr;
  }
}

}  // namespace util
}  // namespace rsb

using namespace rsb;

extern "C" {

int rsb_source_margin(const double* img_dev, int size, double threshold, int* margin_dev,
                      void* stream) {
  RSB_ARG(img_dev && margin_dev && size >= 1, "rsb_source_margin: bad arguments");
  const int init = size;
  RSB_CUDA(cudaMemcpyAsync(margin_dev, &init, sizeof(int), cudaMemcpyHostToDevice, (cudaStream_t)stream));
  int bx = (size * size + 255) / 256;
  if (bx > 1024) bx = 1024;
  util::margin_kernel<<<bx, 256, 0, (cudaStream_t)stream>>>(img_dev, size, threshold, margin_dev);
  RSB_LAUNCHED();
  return 0;
}

int rsb_resize_linear(const double* src_dev, int src_size, int src_pitch, size_t src_plane_stride,
                      int pad, int planes, double* dst_dev, int dst_size, void* stream) {
  RSB_ARG(src_dev && dst_dev && src_size >= 1 && dst_size >= 1 && pad >= 0 && planes >= 0 &&
              src_pitch >= src_size,
          "rsb_resize_linear: bad arguments");
  if (planes == 0) return 0;
  RSB_ARG(planes <= 65535, "rsb_resize_linear: at most 65535 planes per call");
  int bx = (dst_size * dst_size + 255) / 256;
  if (bx > 1024) bx = 1024;
  util::resize_kernel<<<dim3(bx, planes), 256, 0, (cudaStream_t)stream>>>(
      src_dev, src_size, src_pitch, pad, dst_dev, dst_size, src_plane_stride,
      (size_t)dst_size * dst_size);
  RSB_LAUNCHED();
  return 0;
}

}  // extern "C"
