// Protoplanetary-disk images (reference src/radiosim/ppdisks/ppdisks.py).
//
// Per image the host packer (radiosim_b200/ppdisks) replays the RNG chain and hands over,
// per component (core + rings): the two ellipses whose XOR is the mask, the smoothing sigma
// and the amplitude.  All work happens on the fixed 900 x 900 grid of the reference:
//
//   mask_kernel   : _create_ellipse (ppdisks.py:188-236) + the ring XOR (:269), bit-exact:
//                   the reference evaluates the membership test in FLOAT32 (int64 tensor x
//                   numpy scalar promotes to torch's default dtype), so the same IEEE
//                   single-precision mul/add/div sequence is issued here without FMA
//                   contraction.
//   hconv_kernel  : _smooth/_convolve2d (:398-451) is a CIRCULAR convolution with a
//   vconv_kernel    sum-normalised Gaussian sampled at half-integer offsets; it is separable:
//                   out[i,j] = sum_kl mask[k,l] h[(i-k+449)%900] h[(j-l+449)%900].  Row pass
//                   from a shared-memory copy of the byte mask, column pass fused with the
//                   scale-accumulate over components and the abs (:151-152); only the crop
//                   the image uses (:154-161) is produced.
//   resize_kernel : torchvision antialiased bilinear resize (:166-171, ATen
//                   _upsample_bilinear2d_aa) of the crop + paste into the 1e-10 canvas
//                   (:178-185), float32 like the reference's output.
#include "common.cuh"

namespace rsb {
namespace ppdisk {

constexpr int kFix = 900;  // ppdisks.py:131
constexpr int kMaxSigma = 16;

struct Comp {      // one smoothed mask
  float ca, sa;    // float32(cos(alpha)), float32(sin(alpha))
  float a2_in, b2_in, a2_out, b2_out;  // float32(a^2), float32(b^2); a2_in <= 0: no inner ellipse
  float scale;     // amplitude exp(m d + n) + noise (ppdisks.py:477-483)
  int sigma;       // smoothing width (integer in the reference)
  int image;       // owning image
  int pad[3];
};

struct Image {
  int comp0, ncomp;  // components of this image: comp0 .. comp0+ncomp-1 (core first)
  int lo, hi;        // crop rows/cols [lo, hi) of the 900 grid (ppdisks.py:158-161)
  int new_size;      // resize target (even)
  int pad[3];
};

__device__ __forceinline__ bool inside(float dr, float dc, float ca, float sa, float a2, float b2) {
  // ((xx-x0) cos + (yy-y0) sin)^2 / a^2 + ((xx-x0) sin - (yy-y0) cos)^2 / b^2 <= 1, float32,
  // every operation rounded on its own (torch CPU kernels, no fused multiply-add)
  float t1 = __fadd_rn(__fmul_rn(dr, ca), __fmul_rn(dc, sa));
  t1 = __fmul_rn(t1, t1);
  float t2 = __fsub_rn(__fmul_rn(dr, sa), __fmul_rn(dc, ca));
  t2 = __fmul_rn(t2, t2);
  const float q = __fadd_rn(__fdiv_rn(t1, a2), __fdiv_rn(t2, b2));
  return q <= 1.0f;
}

// grid: (rows, components); one CTA per mask row
__global__ void __launch_bounds__(256) mask_kernel(const Comp* __restrict__ comps,
                                                   unsigned char* __restrict__ masks) {
  const Comp c = comps[blockIdx.y];
  const int i = blockIdx.x;
  const float dr = (float)(i - kFix / 2);
  unsigned char* row = masks + ((size_t)blockIdx.y * kFix + i) * kFix;
  for (int j = threadIdx.x; j < kFix; j += blockDim.x) {
    const float dc = (float)(j - kFix / 2);
    bool m = inside(dr, dc, c.ca, c.sa, c.a2_out, c.b2_out);
    if (c.a2_in > 0.f) m ^= inside(dr, dc, c.ca, c.sa, c.a2_in, c.b2_in);
    row[j] = m ? 1 : 0;
  }
}

// Row pass: tmp[c][i][j] = sum_l mask[c][i][l] h[(j-l+449)%900]  for crop rows' SOURCE rows.
// grid: (rows, components).  h is truncated at +-kReach(sigma) taps (tails < 1e-14).
__device__ __forceinline__ int reach(int sigma) { return 8 * sigma + 1; }

__global__ void __launch_bounds__(256) hconv_kernel(const Comp* __restrict__ comps,
                                                    const unsigned char* __restrict__ masks,
                                                    const float* __restrict__ htab,  // [kMaxSigma+1][900]
                                                    float* __restrict__ tmp) {
  __shared__ unsigned char mrow[kFix];
  __shared__ int any;
  const Comp c = comps[blockIdx.y];
  const int i = blockIdx.x;
  const unsigned char* src = masks + ((size_t)blockIdx.y * kFix + i) * kFix;
  if (threadIdx.x == 0) any = 0;
  __syncthreads();
  int mine = 0;
  for (int j = threadIdx.x; j < kFix; j += blockDim.x) {
    const unsigned char m = src[j];
    mrow[j] = m;
    mine |= m;
  }
  if (mine) any = 1;
  __syncthreads();
  float* dst = tmp + ((size_t)blockIdx.y * kFix + i) * kFix;
  if (!any) {  // empty mask row: zeros
    for (int j = threadIdx.x; j < kFix; j += blockDim.x) dst[j] = 0.f;
    return;
  }
  const float* h = htab + (size_t)c.sigma * kFix;
  const int R = reach(c.sigma);
  for (int j = threadIdx.x; j < kFix; j += blockDim.x) {
    // taps t = (j - l + 449) % 900 within [449.5 - R, 449.5 + R]  <=>  l = j - d, d = t - 449
    float s = 0.f;
    for (int d = -R + 1; d <= R; ++d) {
      int l = j - d;
      l += (l < 0) ? kFix : 0;
      l -= (l >= kFix) ? kFix : 0;
      s = fmaf((float)mrow[l], h[d + 449], s);
    }
    dst[j] = s;
  }
}

// Column pass over the crop + scale-accumulate + abs.
// grid: (crop column blocks of 128, crop rows, images)
__global__ void __launch_bounds__(128) vconv_kernel(const Comp* __restrict__ comps,
                                                    const Image* __restrict__ images,
                                                    const float* __restrict__ htab,
                                                    const float* __restrict__ tmp,
                                                    float* __restrict__ acc) {
  const Image im = images[blockIdx.z];
  const int i = im.lo + blockIdx.y;
  const int j = im.lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= im.hi || j >= im.hi) return;
  float core = 0.f, rings = 0.f;
  for (int q = 0; q < im.ncomp; ++q) {
    const Comp c = comps[im.comp0 + q];
    const float* h = htab + (size_t)c.sigma * kFix;
    const float* t = tmp + (size_t)(im.comp0 + q) * kFix * kFix;
    const int R = reach(c.sigma);
    float s = 0.f;
    for (int d = -R + 1; d <= R; ++d) {
      int k = i - d;
      k += (k < 0) ? kFix : 0;
      k -= (k >= kFix) ? kFix : 0;
      s = fmaf(t[(size_t)k * kFix + j], h[d + 449], s);
    }
    // core in "float64" position, rings summed one after the other into a float32 accumulator
    // (els_full of ppdisks.py:139-149)
    if (q == 0) core = s * c.scale;
    else rings = rings + s * c.scale;
  }
  acc[((size_t)blockIdx.z * kFix + i) * kFix + j] = fabsf(core + rings);
}

// Antialiased bilinear resize of acc[lo:hi, lo:hi] to new_size^2 + paste into the canvas.
// grid: (canvas column blocks, canvas rows, images); every canvas pixel written once.
__device__ __forceinline__ void aa_window(int d, double scale, double support, int S, int* xmin,
                                          int* xsize, double* center) {
  *center = scale * (d + 0.5);
  int lo = (int)(*center - support + 0.5);
  lo = lo < 0 ? 0 : lo;
  int hi = (int)(*center + support + 0.5);
  hi = hi > S ? S : hi;
  *xmin = lo;
  *xsize = hi - lo;
}

__global__ void __launch_bounds__(128) resize_kernel(const Image* __restrict__ images,
                                                     const float* __restrict__ acc, int N,
                                                     float* __restrict__ out) {
  const Image im = images[blockIdx.z];
  const int r = blockIdx.y, c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  const float bg = 1e-10f;  // torch.ones(...) * 1e-10 in float32
  float v = bg;
  const int half = im.new_size / 2, c0 = N / 2 - half;
  const int dr = r - c0, dc = c - c0;
  if (dr >= 0 && dr < im.new_size && dc >= 0 && dc < im.new_size) {
    const int S = im.hi - im.lo;
    const double scale = (double)S / (double)im.new_size;
    const double support = scale >= 1.0 ? scale : 1.0;
    const double invscale = scale >= 1.0 ? 1.0 / scale : 1.0;
    int rmin, rsize, cmin, csize;
    double rcen, ccen;
    aa_window(dr, scale, support, S, &rmin, &rsize, &rcen);
    aa_window(dc, scale, support, S, &cmin, &csize, &ccen);
    double wr_tot = 0.0, wc_tot = 0.0;
    for (int a = 0; a < rsize; ++a) wr_tot += fmax(0.0, 1.0 - fabs((a + rmin - rcen + 0.5) * invscale));
    for (int b = 0; b < csize; ++b) wc_tot += fmax(0.0, 1.0 - fabs((b + cmin - ccen + 0.5) * invscale));
    const float* src = acc + ((size_t)blockIdx.z * kFix + im.lo) * kFix + im.lo;
    double s = 0.0;
    for (int a = 0; a < rsize; ++a) {
      const double wr = fmax(0.0, 1.0 - fabs((a + rmin - rcen + 0.5) * invscale));
      if (wr == 0.0) continue;
      double row = 0.0;
      const float* p = src + (size_t)(a + rmin) * kFix + cmin;
      for (int b = 0; b < csize; ++b) {
        const double wc = fmax(0.0, 1.0 - fabs((b + cmin - ccen + 0.5) * invscale));
        row += wc * (double)p[b];
      }
      s += wr * row;
    }
    s /= (wr_tot * wc_tot);
    v = (float)((double)bg + s);  // float32 canvas += float64 patch (ppdisks.py:178-183)
  }
  out[((size_t)blockIdx.z * N + r) * N + c] = v;
}

}  // namespace ppdisk
}  // namespace rsb

using namespace rsb;
using namespace rsb::ppdisk;

extern "C" {

int rsb_ppdisk_workspace_bytes(int n_images, int n_comps, size_t* bytes) {
  RSB_ARG(bytes && n_images >= 0 && n_comps >= 0, "rsb_ppdisk_workspace_bytes: bad arguments");
  const size_t px = (size_t)kFix * kFix;
  size_t b = ((size_t)n_comps * px + 255) & ~(size_t)255;  // byte masks
  b += (size_t)n_comps * px * sizeof(float);               // row-pass result
  b += (size_t)n_images * px * sizeof(float);              // accumulated 900 x 900 image
  *bytes = b + 256;
  return 0;
}

int rsb_ppdisk_masks(const void* comps, int n_comps, unsigned char* masks_dev, void* stream) {
  RSB_ARG(n_comps >= 0, "rsb_ppdisk_masks: n_comps < 0");
  if (n_comps == 0) return 0;
  RSB_ARG(comps && masks_dev, "rsb_ppdisk_masks: NULL buffer");
  for (int c0 = 0; c0 < n_comps; c0 += 65535) {
    const int cnt = (n_comps - c0) < 65535 ? (n_comps - c0) : 65535;
    mask_kernel<<<dim3(kFix, cnt), 256, 0, (cudaStream_t)stream>>>(
        (const Comp*)comps + c0, masks_dev + (size_t)c0 * kFix * kFix);
    RSB_LAUNCHED();
  }
  return 0;
}

int rsb_ppdisk_render(const void* comps, int n_comps, const void* images, int n_images,
                      const float* htab_dev, int img_size, float* out_dev, void* workspace,
                      size_t workspace_bytes, void* stream) {
  RSB_ARG(n_images >= 0 && n_comps >= 0 && img_size >= 2, "rsb_ppdisk_render: bad sizes");
  if (n_images == 0) return 0;
  RSB_ARG(n_comps <= 65535 && n_images <= 65535, "rsb_ppdisk_render: at most 65535 components/images per call");
  RSB_ARG(comps && images && htab_dev && out_dev && workspace, "rsb_ppdisk_render: NULL buffer");
  size_t need = 0;
  if (int e = rsb_ppdisk_workspace_bytes(n_images, n_comps, &need)) return e;
  RSB_ARG(workspace_bytes >= need, "rsb_ppdisk_render: workspace too small (%zu < %zu bytes)",
          workspace_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t px = (size_t)kFix * kFix;
  unsigned char* masks = (unsigned char*)workspace;
  float* tmp = (float*)((char*)workspace + (((size_t)n_comps * px + 255) & ~(size_t)255));
  float* acc = tmp + (size_t)n_comps * px;
  if (int e = rsb_ppdisk_masks(comps, n_comps, masks, stream)) return e;
  hconv_kernel<<<dim3(kFix, n_comps), 256, 0, st>>>((const Comp*)comps, masks, htab_dev, tmp);
  RSB_LAUNCHED();
  vconv_kernel<<<dim3((kFix + 127) / 128, kFix, n_images), 128, 0, st>>>(
      (const Comp*)comps, (const Image*)images, htab_dev, tmp, acc);
  RSB_LAUNCHED();
  resize_kernel<<<dim3((img_size + 127) / 128, img_size, n_images), 128, 0, st>>>(
      (const Image*)images, acc, img_size, out_dev);
  RSB_LAUNCHED();
  return 0;
}

}  // extern "C"
