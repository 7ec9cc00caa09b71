// MOJAVE-like sources (reference src/radiosim/mojave/mojave.py).
//
// The host packer replays each child generator's draws and sends, per image, a recipe:
//   * up to two "jets", each a sum of skewed-Gaussian blobs  amp * normal[row] * skew[col]
//     (src/radiosim/utils/gauss.py:100-150; the two max-normalised 1-D profiles are
//     evaluated on the host with the reference's scipy calls and shipped as vectors),
//     zeroed where <= 1e-8 (np.isclose, gauss.py:148), and a swirl (mojave.py:222-261);
//   * up to three rotated Gaussians (gauss.py:50-97) for the core / compact components;
//   * the final rotation angle (mojave.py:346,463,626) and, for compact sources, the
//     recentre + rescale step (mojave.py:628-635).
//
//   pre_kernel    : G = swirl(jet_r) + flip(swirl(jet_l)) + gaussians, every pixel evaluated
//                   directly from the recipe (the swirl's bilinear taps re-evaluate the blob
//                   sum at 4 integer pixels; nothing but G is stored).
//   rot_kernel    : skimage.transform.rotate = bilinear gather of G around (N/2-0.5, N/2-0.5),
//                   zero outside, isclose zeroing; compact sources also reduce (max, first
//                   arg-max) and finish in fin_kernel (roll + /max * amp).
// Coordinates of both warps are computed in float64 (a 1e-7 rad error on a 100 px radius
// would already cost 1e-5 px), pixel values in float32.
#include "common.cuh"

namespace rsb {
namespace mojave {

struct Gauss {  // twodgaussian parameters (float64) -- amp == 0: unused
  double amp, x, y, sx, sy, rot;
};

struct Recipe {
  Gauss g[3];
  double strength[2], radius[2];  // swirl of jet 0 / jet 1 (centre is (N/2, N/2))
  double angle;                   // final rotation, degrees
  double amp;                     // compact: final amplitude
  int blob0[2], nblob[2];         // blobs of jet 0 / jet 1 in the profile table
  int flip1;                      // jet 1 is added flipped in both axes (two-jet sources)
  int compact;                    // recentre on the maximum and rescale to amp
  int pad[2];
};

struct Blob {
  float amp;
  int pad;
  long long prof;  // offset (floats) of normal[N] followed by skew[N]
};

__device__ __forceinline__ float blob_sum(const Blob* __restrict__ blobs, int nb,
                                          const float* __restrict__ prof, int N, int r, int c) {
  float s = 0.f;
  for (int b = 0; b < nb; ++b) {
    const float* p = prof + blobs[b].prof;
    // (normal * skew) * amp, then np.isclose(.., 0) -> 0
    const float v = (p[r] * p[N + c]) * blobs[b].amp;
    s += (fabsf(v) <= 1e-8f) ? 0.f : v;
  }
  return s;
}

__device__ __forceinline__ double mirror(double x, int n) {
  // ndimage 'mirror': reflect about the centres of the edge pixels, period 2(n-1)
  if (n == 1) return 0.0;
  const double period = 2.0 * (n - 1);
  x = fabs(x);
  x -= period * floor(x / period);
  return x > (n - 1) ? period - x : x;
}

// swirl(jet)(y, x): inverse map + bilinear (skimage.transform.swirl, order 1, mode reflect)
__device__ float swirled(const Blob* blobs, int nb, const float* prof, int N, double strength,
                         double radius, int y, int x) {
  if (nb == 0) return 0.f;
  const double c0 = N / 2.0;
  const double dx = x - c0, dy = y - c0;
  const double rho = sqrt(dx * dx + dy * dy);
  const double rr = radius / 5.0 * 0.6931471805599453;
  // (x0 + rho cos(theta0 + phi), y0 + rho sin(theta0 + phi)) with theta0 = atan2(dy, dx) is the
  // offset (dx, dy) turned by phi = strength exp(-rho / r'): no atan2, no second rho * cos / sin
  double sp, cp;
  sincos(strength * exp(-rho / rr), &sp, &cp);
  const double sxp = mirror(c0 + (dx * cp - dy * sp), N), syp = mirror(c0 + (dx * sp + dy * cp), N);
  const int x0 = (int)floor(sxp), y0 = (int)floor(syp);
  const float fx = (float)(sxp - x0), fy = (float)(syp - y0);
  const int x1 = min(x0 + 1, N - 1), y1 = min(y0 + 1, N - 1);
  const float v00 = blob_sum(blobs, nb, prof, N, y0, x0), v01 = blob_sum(blobs, nb, prof, N, y0, x1);
  const float v10 = blob_sum(blobs, nb, prof, N, y1, x0), v11 = blob_sum(blobs, nb, prof, N, y1, x1);
  const float top = fmaf(fx, v01 - v00, v00), bot = fmaf(fx, v11 - v10, v10);
  return fmaf(fy, bot - top, top);
}

__device__ __forceinline__ float gauss_px(const Gauss& g, int r, int c) {
  if (g.amp == 0.0) return 0.f;
  // operation order of gauss.py:77-88 in float64 (3 per image: negligible)
  const double cr = cos(g.rot), sr = sin(g.rot);
  const double ra = g.y * cr - g.x * sr, rb = g.y * sr + g.x * cr;
  const double ua = r * cr - c * sr, ub = r * sr + c * cr;
  const double a = (ra - ua) / g.sy, b = (rb - ub) / g.sx;
  return (float)(g.amp * exp(-(a * a + b * b) / 2.0));
}

// grid: (pixel blocks, images)
__global__ void __launch_bounds__(256) pre_kernel(const Recipe* __restrict__ recipes,
                                                  const Blob* __restrict__ blobs,
                                                  const float* __restrict__ prof, int N,
                                                  float* __restrict__ G) {
  const Recipe& R = recipes[blockIdx.y];
  float* g = G + (size_t)blockIdx.y * N * N;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < N * N; e += gridDim.x * blockDim.x) {
    const int y = e / N, x = e - y * N;
    float v = swirled(blobs + R.blob0[0], R.nblob[0], prof, N, R.strength[0], R.radius[0], y, x);
    if (R.nblob[1] > 0) {
      const int yy = R.flip1 ? N - 1 - y : y, xx = R.flip1 ? N - 1 - x : x;
      v += swirled(blobs + R.blob0[1], R.nblob[1], prof, N, R.strength[1], R.radius[1], yy, xx);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) v += gauss_px(R.g[k], y, x);
    g[e] = v;
  }
}

// ---- blobs as drawn -> max-normalised 1-D profiles + amplitudes ---------------------------
// skewed_gauss (gauss.py:131-148): normal = norm.pdf(rows; y, width) / max, skew =
// skewnorm.pdf(cols; a, x, length) / max, maxima over the INTEGER samples.  A shock's amplitude is
// the value of the jet built so far at one pixel times a drawn factor (mojave.py:190), so the blobs
// of a jet are finished one after the other.  Float64 throughout; one CTA per jet.
struct BlobParam {  // = rsb::mjpack::BlobParam (csrc/mojave_pack.cu)
  double x, y, amp, width, length, a;
  int ref_r, ref_c;
  int pad[2];
};
constexpr int kMaxJetBlobs = 32;

__device__ __forceinline__ double normal_pdf_at(const BlobParam& p, int i) {
  const double t = ((double)i - p.y) / p.width;
  return exp(-(t * t) / 2.0) / 2.5066282746310002 / p.width;   // sqrt(2 pi)
}
__device__ __forceinline__ double skew_pdf_at(const BlobParam& p, int i) {
  const double u = ((double)i - p.x) / p.length;
  // 2 phi(u) Phi(a u) / length, Phi(z) = erfc(-z / sqrt 2) / 2
  return 2.0 * (exp(-(u * u) / 2.0) / 2.5066282746310002) * (0.5 * erfc(-(p.a * u) * 0.7071067811865476)) / p.length;
}

__global__ void __launch_bounds__(256) blob_kernel(const Recipe* __restrict__ recipes,
                                                   const BlobParam* __restrict__ bp, int N,
                                                   Blob* __restrict__ blobs, float* __restrict__ prof) {
  const Recipe& R = recipes[blockIdx.x >> 1];
  const int j = blockIdx.x & 1, b0 = R.blob0[j], nb = min(R.nblob[j], kMaxJetBlobs);
  if (nb <= 0) return;
  __shared__ double s_nmax[kMaxJetBlobs], s_smax[kMaxJetBlobs], s_amp[kMaxJetBlobs];
  __shared__ double s_red[2][8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int k = 0; k < nb; ++k) {
    const BlobParam p = bp[b0 + k];
    double nm = 0.0, sm = 0.0;   // pdf values are >= 0
    for (int i = tid; i < N; i += blockDim.x) {
      nm = fmax(nm, normal_pdf_at(p, i));
      sm = fmax(sm, skew_pdf_at(p, i));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      nm = fmax(nm, __shfl_xor_sync(0xffffffffu, nm, o));
      sm = fmax(sm, __shfl_xor_sync(0xffffffffu, sm, o));
    }
    if (lane == 0) { s_red[0][warp] = nm; s_red[1][warp] = sm; }
    __syncthreads();
    if (tid == 0) {
      double a = 0.0, b = 0.0;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) { a = fmax(a, s_red[0][w]); b = fmax(b, s_red[1][w]); }
      s_nmax[k] = a;
      s_smax[k] = b;
      double amp = p.amp;
      if (p.ref_r >= 0) {  // jet.pixel(r, c) of the blobs before this one, summed in build order
        double sum = 0.0;
        for (int q = 0; q < k; ++q) {
          const BlobParam pq = bp[b0 + q];
          const double v = ((normal_pdf_at(pq, p.ref_r) / s_nmax[q]) * (skew_pdf_at(pq, p.ref_c) / s_smax[q])) * s_amp[q];
          sum = sum + (fabs(v) <= 1e-8 ? 0.0 : v);
        }
        amp = sum * p.amp;
      }
      s_amp[k] = amp;
      Blob out;
      out.amp = (float)amp; out.pad = 0; out.prof = (long long)(b0 + k) * 2 * N;
      blobs[b0 + k] = out;
    }
    __syncthreads();
    const double inm = s_nmax[k], ism = s_smax[k];
    float* dst = prof + (size_t)(b0 + k) * 2 * N;
    for (int i = tid; i < N; i += blockDim.x) {
      dst[i] = (float)(normal_pdf_at(p, i) / inm);
      dst[N + i] = (float)(skew_pdf_at(p, i) / ism);
    }
  }
}

template <typename OutT>
__global__ void __launch_bounds__(256) rot_kernel(const Recipe* __restrict__ recipes,
                                                  const float* __restrict__ G, int N,
                                                  OutT* __restrict__ out, float* __restrict__ rot_tmp,
                                                  unsigned long long* __restrict__ best) {
  const Recipe& R = recipes[blockIdx.y];
  const float* g = G + (size_t)blockIdx.y * N * N;
  const double ang = R.angle * 0.017453292519943295;
  double sa, ca;
  sincos(ang, &sa, &ca);
  const double cx = N / 2.0 - 0.5, cy = N / 2.0 - 0.5;
  const double m02 = cx - ca * cx + sa * cy, m12 = cy - sa * cx - ca * cy;
  unsigned long long my_best = 0ull;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < N * N; e += gridDim.x * blockDim.x) {
    const int y = e / N, x = e - y * N;
    const double sx = ca * x - sa * y + m02, sy = sa * x + ca * y + m12;
    const double fxd = floor(sx), fyd = floor(sy);
    const int x0 = (int)fxd, y0 = (int)fyd;
    const float fx = (float)(sx - fxd), fy = (float)(sy - fyd);
    auto at = [&](int yy, int xx) -> float {
      return (yy >= 0 && yy < N && xx >= 0 && xx < N) ? g[(size_t)yy * N + xx] : 0.f;
    };
    const float v00 = at(y0, x0), v01 = at(y0, x0 + 1), v10 = at(y0 + 1, x0), v11 = at(y0 + 1, x0 + 1);
    const float top = fmaf(fx, v01 - v00, v00), bot = fmaf(fx, v11 - v10, v10);
    float v = fmaf(fy, bot - top, top);
    if (R.compact) {
      rot_tmp[(size_t)blockIdx.y * N * N + e] = v;  // zeroing happens after the roll
      const unsigned long long key =
          ((unsigned long long)__float_as_uint(fmaxf(v, 0.f)) << 32) | (unsigned)(~(unsigned)e);
      my_best = key > my_best ? key : my_best;
    } else {
      if (fabsf(v) <= 1e-8f) v = 0.f;  // glx[np.isclose(glx, 0)] = 0
      out[(size_t)blockIdx.y * N * N + e] = (OutT)v;
    }
  }
  if (R.compact) {
    my_best = warp_max_u64(my_best);
    if ((threadIdx.x & 31) == 0) atomicMax(best + blockIdx.y, my_best);
  }
}

template <typename OutT>
__global__ void __launch_bounds__(256) fin_kernel(const Recipe* __restrict__ recipes,
                                                  const float* __restrict__ rot_tmp, int N,
                                                  const unsigned long long* __restrict__ best,
                                                  OutT* __restrict__ out) {
  const Recipe& R = recipes[blockIdx.y];
  if (!R.compact) return;
  const unsigned long long b = best[blockIdx.y];
  const float peak = __uint_as_float((unsigned)(b >> 32));
  const unsigned idx = ~(unsigned)(b & 0xffffffffull);
  const int py = (int)(idx / (unsigned)N), px = (int)(idx % (unsigned)N);
  const int shy = N / 2 - py, shx = N / 2 - px;  // np.roll(glx, shift, axis=[0, 1])
  const float scale = (float)(R.amp / (double)peak);
  const float* src = rot_tmp + (size_t)blockIdx.y * N * N;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < N * N; e += gridDim.x * blockDim.x) {
    const int y = e / N, x = e - y * N;
    int sy = (y - shy) % N, sx = (x - shx) % N;
    sy += sy < 0 ? N : 0;
    sx += sx < 0 ? N : 0;
    float v = src[(size_t)sy * N + sx];
    if (fabsf(v) <= 1e-8f) v = 0.f;
    out[(size_t)blockIdx.y * N * N + e] = (OutT)(v * scale);
  }
}

// skimage.transform.swirl of an ARBITRARY float64 image (add_swirl as a stand-alone drop-in; the bundle
// path swirls analytically inside pre_kernel).  Same restatement as `swirled` above: inverse map, the
// offset vector turned by strength * exp(-rho / r'), ndimage 'mirror' boundary, order-1 interpolation in
// float64, result clipped to the input's range [lo, hi].  grid: (column blocks, rows)
__global__ void __launch_bounds__(256) swirl_image_kernel(const double* __restrict__ in, int rows, int cols,
                                                          double cx, double cy, double strength, double radius,
                                                          double lo, double hi, double* __restrict__ out) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= cols) return;
  const double dx = x - cx, dy = y - cy;
  const double rho = sqrt(dx * dx + dy * dy);
  const double rr = radius / 5.0 * 0.6931471805599453;
  double sp, cp;
  sincos(strength * exp(-rho / rr), &sp, &cp);
  const double sxp = mirror(cx + (dx * cp - dy * sp), cols), syp = mirror(cy + (dx * sp + dy * cp), rows);
  const int x0 = (int)floor(sxp), y0 = (int)floor(syp);
  const double fx = sxp - x0, fy = syp - y0;
  const int x1 = min(x0 + 1, cols - 1), y1 = min(y0 + 1, rows - 1);  // (weight 0 when clamped)
  const double v00 = in[(size_t)y0 * cols + x0], v01 = in[(size_t)y0 * cols + x1];
  const double v10 = in[(size_t)y1 * cols + x0], v11 = in[(size_t)y1 * cols + x1];
  const double v = (1.0 - fy) * ((1.0 - fx) * v00 + fx * v01) + fy * ((1.0 - fx) * v10 + fx * v11);
  out[(size_t)y * cols + x] = fmin(fmax(v, lo), hi);
}

}  // namespace mojave
}  // namespace rsb

using namespace rsb;
using namespace rsb::mojave;

extern "C" {

int rsb_mojave_workspace_bytes(int n, int size, size_t* bytes) {
  RSB_ARG(bytes && n >= 0 && size >= 2, "rsb_mojave_workspace_bytes: bad arguments");
  const size_t px = (size_t)size * size;
  *bytes = 2 * (size_t)n * px * sizeof(float) + (((size_t)n * 8 + 255) & ~(size_t)255) + 256;
  return 0;
}

int rsb_mojave_blobs(const void* recipes, int n, const void* blob_params, int n_blobs, int size,
                     void* blobs_out, float* profiles_out, void* stream) {
  RSB_ARG(n >= 0 && n_blobs >= 0 && size >= 2, "rsb_mojave_blobs: bad n / n_blobs / size (%d, %d, %d)", n, n_blobs, size);
  if (n == 0 || n_blobs == 0) return 0;
  RSB_ARG(recipes && blob_params && blobs_out && profiles_out, "rsb_mojave_blobs: NULL buffer");
  blob_kernel<<<2 * n, 256, 0, (cudaStream_t)stream>>>((const Recipe*)recipes, (const BlobParam*)blob_params, size,
                                                     (Blob*)blobs_out, profiles_out);
  RSB_LAUNCHED();
  return 0;
}

int rsb_swirl(const double* image_dev, int rows, int cols, double center_x, double center_y, double strength,
              double radius, double clip_lo, double clip_hi, double* out_dev, void* stream) {
  RSB_ARG(rows >= 1 && cols >= 1 && rows <= 65535, "rsb_swirl: bad shape (%d, %d)", rows, cols);
  RSB_ARG(image_dev && out_dev && image_dev != out_dev, "rsb_swirl: NULL buffer or in-place call");
  RSB_ARG(radius != 0.0, "rsb_swirl: radius must not be zero");
  swirl_image_kernel<<<dim3((cols + 255) / 256, rows), 256, 0, (cudaStream_t)stream>>>(
      image_dev, rows, cols, center_x, center_y, strength, radius, clip_lo, clip_hi, out_dev);
  RSB_LAUNCHED();
  return 0;
}

int rsb_mojave_render(const void* recipes, int n, const void* blobs, const float* profiles,
                      int size, unsigned flags, void* out, void* workspace, size_t workspace_bytes,
                      void* stream) {
  RSB_ARG(n >= 0 && size >= 2 && size <= 16384, "rsb_mojave_render: bad n/size (%d, %d)", n, size);
  if (n == 0) return 0;
  RSB_ARG(n <= 65535, "rsb_mojave_render: at most 65535 images per call");
  RSB_ARG(recipes && blobs && profiles && out && workspace, "rsb_mojave_render: NULL buffer");
  size_t need = 0;
  if (int e = rsb_mojave_workspace_bytes(n, size, &need)) return e;
  RSB_ARG(workspace_bytes >= need, "rsb_mojave_render: workspace too small (%zu < %zu bytes)",
          workspace_bytes, need);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t px = (size_t)size * size;
  float* G = (float*)workspace;
  float* rot_tmp = G + (size_t)n * px;
  unsigned long long* best = (unsigned long long*)(rot_tmp + (size_t)n * px);
  best = (unsigned long long*)(((uintptr_t)best + 255) & ~(uintptr_t)255);
  RSB_CUDA(cudaMemsetAsync(best, 0, (size_t)n * 8, st));
  int bx = (int)((px + 255) / 256);
  if (bx > 256) bx = 256;
  const dim3 grid(bx, n);
  pre_kernel<<<grid, 256, 0, st>>>((const Recipe*)recipes, (const Blob*)blobs, profiles, size, G);
  RSB_LAUNCHED();
  if (flags & RSB_JET_OUT_F64) {
    rot_kernel<double><<<grid, 256, 0, st>>>((const Recipe*)recipes, G, size, (double*)out, rot_tmp, best);
    RSB_LAUNCHED();
    fin_kernel<double><<<grid, 256, 0, st>>>((const Recipe*)recipes, rot_tmp, size, best, (double*)out);
    RSB_LAUNCHED();
  } else {
    rot_kernel<float><<<grid, 256, 0, st>>>((const Recipe*)recipes, G, size, (float*)out, rot_tmp, best);
    RSB_LAUNCHED();
    fin_kernel<float><<<grid, 256, 0, st>>>((const Recipe*)recipes, rot_tmp, size, best, (float*)out);
    RSB_LAUNCHED();
  }
  return 0;
}

}  // extern "C"
