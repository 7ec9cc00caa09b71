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
//   hrun_kernel   : _smooth/_convolve2d (:398-451) is a CIRCULAR convolution with a
//   vconv_kernel    sum-normalised Gaussian sampled at half-integer offsets; it is separable:
//                   out[i,j] = sum_kl mask[k,l] h[(i-k+449)%900] h[(j-l+449)%900].  Row pass
//                   from the RUNS of each mask row and cumulative taps (O(transitions) per
//                   pixel), column pass from shared-memory tiles (4 x 4 outputs per thread)
//                   fused with the scale-accumulate over components and the abs (:151-152);
//                   only the crop the image uses (:154-161) is produced.
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

// Row pass: tmp[c][i][j] = sum_l mask[c][i][l] h[(j-l+449)%900], taps truncated at +-reach(sigma)
// (tails < 1e-14).  A mask row is a handful of RUNS (an ellipse row is one interval, a ring row
// two), so with the cumulative taps G[d] = sum_{u <= d} h[u + 449] a run [l0, l1] contributes
// G[j - l0] - G[j - l1 - 1]: the pass costs O(transitions) per pixel instead of O(taps), and is
// summed in float64.  The three terms per transition are the circular wrap (j - t, j - t +- 900).
__device__ __forceinline__ int reach(int sigma) { return 8 * sigma + 1; }
constexpr int kMaxReach = 8 * kMaxSigma + 1;
constexpr int kMaxTrans = 64;  // transitions per mask row kept as a list (more: plain summation)

// gcum[s][k] = sum_{d = -R+1}^{k - R} h_s[d + 449], k = 0 .. 2R (gcum[s][0] = 0), R = reach(s)
__global__ void gcum_kernel(const float* __restrict__ htab, double* __restrict__ gcum) {
  const int s = blockIdx.x;
  if (s < 1 || threadIdx.x != 0) return;
  const int R = reach(s);
  double acc = 0.0;
  double* g = gcum + (size_t)s * (2 * kMaxReach + 1);
  g[0] = 0.0;
  for (int k = 1; k <= 2 * R; ++k) {
    acc += (double)htab[(size_t)s * kFix + (k - R) + 449];
    g[k] = acc;
  }
}

// grid: (rows, components); block 128.  Only the crop's columns [lo, hi) are produced, and only the
// rows the column pass reads (crop rows +- reach, circular).  The mask row is evaluated here (same
// float32 sequence as mask_kernel: bit-identical) instead of being read back from memory.
__global__ void __launch_bounds__(128) hrun_kernel(const Comp* __restrict__ comps,
                                                   const Image* __restrict__ images,
                                                   const double* __restrict__ gcum,
                                                   float* __restrict__ tmp) {
  __shared__ double G[2 * kMaxReach + 1];
  __shared__ unsigned char mrow[kFix + 4];
  __shared__ int tlist[4][kMaxTrans / 4];   // transitions found by each warp in its quarter of the row
  __shared__ int tcount[4];
  __shared__ int tpos[kMaxTrans];
  const Comp c = comps[blockIdx.y];
  const Image im = images[c.image];
  const int i = blockIdx.x, R = reach(c.sigma), tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  {
    int off = i - (im.lo - R);   // position of this row among the rows the column pass reads
    off += (off < 0) ? kFix : 0;
    off -= (off >= kFix) ? kFix : 0;
    if (off >= (im.hi - im.lo) + 2 * R && (im.hi - im.lo) + 2 * R < kFix) return;
  }
  for (int k = tid; k <= 2 * R; k += blockDim.x) G[k] = gcum[(size_t)c.sigma * (2 * kMaxReach + 1) + k];
  const float dr = (float)(i - kFix / 2);
  for (int j = tid; j < kFix; j += blockDim.x) {
    const float dc = (float)(j - kFix / 2);
    bool m = inside(dr, dc, c.ca, c.sa, c.a2_out, c.b2_out);
    if (c.a2_in > 0.f) m ^= inside(dr, dc, c.ca, c.sa, c.a2_in, c.b2_in);
    mrow[j] = m ? 1 : 0;
  }
  __syncthreads();
  {
    // ordered list of the positions t in [0, 900] where the row changes value: each warp scans a
    // contiguous quarter (the last one also position 900), the lists are concatenated afterwards
    const int q0 = warp * 225, q1 = (warp == 3) ? kFix + 1 : q0 + 225;
    int n = 0;
    for (int j0 = q0; j0 < q1; j0 += 32) {
      const int j = j0 + lane;
      const int cur = (j < kFix) ? mrow[j] : 0, prev = (j >= 1 && j <= kFix) ? mrow[j - 1] : 0;
      const bool tr = (j < q1) && (cur != prev);
      const unsigned m = __ballot_sync(0xffffffffu, tr);
      if (tr) {
        const int pos = n + __popc(m & ((1u << lane) - 1u));
        if (pos < kMaxTrans / 4) tlist[warp][pos] = j;
      }
      n += __popc(m);
    }
    if (lane == 0) tcount[warp] = n;
  }
  __syncthreads();
  const int n0 = tcount[0], n1 = tcount[1], n2 = tcount[2], n3 = tcount[3];
  const bool listed = n0 <= kMaxTrans / 4 && n1 <= kMaxTrans / 4 && n2 <= kMaxTrans / 4 && n3 <= kMaxTrans / 4;
  const int n = n0 + n1 + n2 + n3;
  if (listed && tid < n) {   // (starts and ends alternate: even entries are starts)
    const int w = (tid >= n0) + (tid >= n0 + n1) + (tid >= n0 + n1 + n2);
    const int base = (w > 0 ? n0 : 0) + (w > 1 ? n1 : 0) + (w > 2 ? n2 : 0);
    tpos[tid] = tlist[w][tid - base];
  }
  __syncthreads();
  float* dst = tmp + ((size_t)blockIdx.y * kFix + i) * kFix;
  const double total = G[2 * R];
  if (listed) {
    for (int j = im.lo + tid; j < im.hi; j += blockDim.x) {
      double s = 0.0;
      for (int q = 0; q < n; ++q) {
        const int x = j - tpos[q];   // G_at(x) = sum_{d <= x} g[d] = G[x + R]
        double v = 0.0;
#pragma unroll
        for (int w = -1; w <= 1; ++w) {
          const int k = x + w * kFix + R;          // g defined for d in (-R, R]
          v += (k <= 0) ? 0.0 : (k >= 2 * R ? total : G[k]);
        }
        s += (q & 1) ? -v : v;
      }
      dst[j] = (float)s;
    }
  } else {  // a ragged row (never produced by ellipses): plain summation over the taps
    for (int j = im.lo + tid; j < im.hi; j += blockDim.x) {
      double s = 0.0;
      for (int d = -R + 1; d <= R; ++d) {
        int l = j - d;
        l += (l < 0) ? kFix : 0;
        l -= (l >= kFix) ? kFix : 0;
        if (mrow[l]) s += G[d + R] - G[d + R - 1];
      }
      dst[j] = (float)s;
    }
  }
}

// Column pass over the crop + scale-accumulate + abs:
//   acc[i][j] = | sum_q scale_q sum_d tmp_q[(i - d) % 900][j] h_q[d + 449] |
// One CTA per 64 x 64 output tile, 4 x 4 outputs per thread.  Per component the tile's input rows
// (64 + 2R - 1 of them, circular in the 900 grid) are staged in shared memory; each thread slides down
// its four columns: one 128-bit load of a staged row and one tap feed sixteen FMAs.
constexpr int kVT = 64;                                   // tile edge
constexpr int kVRows = kVT + 2 * kMaxReach;               // staged rows (upper bound)
__global__ void __launch_bounds__(256) vconv_kernel(const Comp* __restrict__ comps,
                                                    const Image* __restrict__ images,
                                                    const float* __restrict__ htab,
                                                    const float* __restrict__ tmp,
                                                    float* __restrict__ acc) {
  extern __shared__ __align__(16) float vsm[];
  float* tile = vsm;                      // [rows][kVT]
  float* taps = vsm + kVRows * kVT;       // [2R]: taps[k] = h[(k - R + 1) + 449], d = k - R + 1
  const Image im = images[blockIdx.z];
  const int i0 = im.lo + blockIdx.y * kVT, j0 = im.lo + blockIdx.x * kVT;
  if (i0 >= im.hi || j0 >= im.hi) return;
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int rl0 = ty * 4, cl0 = tx * 4;
  float core[4][4], rings[4][4];
#pragma unroll
  for (int a = 0; a < 4; ++a)
#pragma unroll
    for (int b = 0; b < 4; ++b) core[a][b] = rings[a][b] = 0.f;
  for (int q = 0; q < im.ncomp; ++q) {
    const Comp c = comps[im.comp0 + q];
    const int R = reach(c.sigma), nrows = kVT + 2 * R - 1;
    const float* t = tmp + (size_t)(im.comp0 + q) * kFix * kFix;
    __syncthreads();  // the previous component's tile is no longer read
    // staged row p holds grid row (i0 - R + p) mod 900: output row rl uses rows p = rl .. rl + 2R - 1
    for (int e = tid; e < nrows * (kVT / 4); e += blockDim.x) {
      const int p = e / (kVT / 4), c4 = (e - p * (kVT / 4)) * 4;
      int k = i0 - R + p;
      k += (k < 0) ? kFix : 0;
      k -= (k >= kFix) ? kFix : 0;
      const float* srow = t + (size_t)k * kFix + j0 + c4;
      // rows hrun_kernel did not produce (below the last output row's window: they only meet zero
      // taps and unstored outputs) are staged as zeros -- 0 x whatever the workspace held is not 0
      int off = k - (im.lo - R);
      off += (off < 0) ? kFix : 0;
      off -= (off >= kFix) ? kFix : 0;
      const int span = (im.hi - im.lo) + 2 * R;
      const bool have = off < span || span >= kFix;
      float4 v;
      v.x = (have && j0 + c4 + 0 < im.hi) ? srow[0] : 0.f;
      v.y = (have && j0 + c4 + 1 < im.hi) ? srow[1] : 0.f;
      v.z = (have && j0 + c4 + 2 < im.hi) ? srow[2] : 0.f;
      v.w = (have && j0 + c4 + 3 < im.hi) ? srow[3] : 0.f;
      *reinterpret_cast<float4*>(tile + p * kVT + c4) = v;
    }
    for (int k = tid; k < 2 * R; k += blockDim.x) taps[k] = htab[(size_t)c.sigma * kFix + (k - R + 1) + 449];
    __syncthreads();
    // out[rl] = sum_{p = rl}^{rl + 2R - 1} tile[p] * g[d],  d = rl + R - p  ->  taps index k = d + R - 1 = rl + 2R - 1 - p
    float s[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) s[a][b] = 0.f;
    // step p = rl0 + m, m = 0 .. 2R + 2: row a (rl = rl0 + a) takes tap index ka = a + 2R - 1 - m when 0 <= ka < 2R.
    // Sliding window of four taps w[a] = taps[a + 2R - 1 - m] (zero outside).
    float w0 = 0.f, w1 = 0.f, w2 = 0.f, w3 = 0.f;   // taps for rows 0..3 at the current step
    // at m = 0: ka = a + 2R - 1: only a = 0 is inside
    const float* trow = tile + rl0 * kVT + cl0;
    const int steps = 2 * R + 3;
    for (int m = 0; m < steps; ++m) {
      // shift: row a at step m uses what row a-1 used at step m-1
      w3 = w2; w2 = w1; w1 = w0;
      const int k0 = 2 * R - 1 - m;
      w0 = (k0 >= 0) ? taps[k0] : 0.f;
      const float4 v = *reinterpret_cast<const float4*>(trow + m * kVT);
      s[0][0] = fmaf(v.x, w0, s[0][0]); s[0][1] = fmaf(v.y, w0, s[0][1]); s[0][2] = fmaf(v.z, w0, s[0][2]); s[0][3] = fmaf(v.w, w0, s[0][3]);
      s[1][0] = fmaf(v.x, w1, s[1][0]); s[1][1] = fmaf(v.y, w1, s[1][1]); s[1][2] = fmaf(v.z, w1, s[1][2]); s[1][3] = fmaf(v.w, w1, s[1][3]);
      s[2][0] = fmaf(v.x, w2, s[2][0]); s[2][1] = fmaf(v.y, w2, s[2][1]); s[2][2] = fmaf(v.z, w2, s[2][2]); s[2][3] = fmaf(v.w, w2, s[2][3]);
      s[3][0] = fmaf(v.x, w3, s[3][0]); s[3][1] = fmaf(v.y, w3, s[3][1]); s[3][2] = fmaf(v.z, w3, s[3][2]); s[3][3] = fmaf(v.w, w3, s[3][3]);
    }
    // core in "float64" position, rings summed one after the other into a float32 accumulator
    // (els_full of ppdisks.py:139-149)
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        if (q == 0) core[a][b] = s[a][b] * c.scale;
        else rings[a][b] = rings[a][b] + s[a][b] * c.scale;
      }
  }
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int i = i0 + rl0 + a;
    if (i >= im.hi) continue;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const int j = j0 + cl0 + b;
      if (j < im.hi) acc[((size_t)blockIdx.z * kFix + i) * kFix + j] = fabsf(core[a][b] + rings[a][b]);
    }
  }
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
  b += (size_t)(kMaxSigma + 1) * (2 * kMaxReach + 1) * sizeof(double) + 256;  // cumulative taps
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
  double* gcum = (double*)(((uintptr_t)(acc + (size_t)n_images * px) + 255) & ~(uintptr_t)255);
  (void)masks;  // (the mask rows are evaluated inside hrun_kernel; rsb_ppdisk_masks serves callers that want them)
  gcum_kernel<<<kMaxSigma + 1, 32, 0, st>>>(htab_dev, gcum);
  RSB_LAUNCHED();
  hrun_kernel<<<dim3(kFix, n_comps), 128, 0, st>>>((const Comp*)comps, (const Image*)images, gcum, tmp);
  RSB_LAUNCHED();
  const size_t vsmem = (size_t)(kVRows * kVT + 2 * kMaxReach) * sizeof(float);
  RSB_CUDA(cudaFuncSetAttribute(vconv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vsmem));
  vconv_kernel<<<dim3((kFix + kVT - 1) / kVT, (kFix + kVT - 1) / kVT, n_images), 256, vsmem, st>>>(
      (const Comp*)comps, (const Image*)images, htab_dev, tmp, acc);
  RSB_LAUNCHED();
  resize_kernel<<<dim3((img_size + 127) / 128, img_size, n_images), 128, 0, st>>>(
      (const Image*)images, acc, img_size, out_dev);
  RSB_LAUNCHED();
  return 0;
}

}  // extern "C"
