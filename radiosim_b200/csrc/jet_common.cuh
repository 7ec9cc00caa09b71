// Pieces shared by the jet kernels (jet.cu: generic/global-scratch path, jet2.cu:
// shared-memory / cluster path).
#pragma once
#include <cmath>

#include "common.cuh"

namespace rsb {
namespace jet {

constexpr int kStrip = 128;  // columns per warp strip: 4 per lane
constexpr int kMaxSlots = RSB_JET_MAX_SLOTS;
constexpr double kEdgeThr = 0.01;  // jet.py:156 / utils.py:90 max_amp

// float32 coefficients of one live component (read warp-uniformly).
//   g(r,c) = amp * 2^(A dc^2 + B dc dr + C dr^2),  dc = c - xc, dr = r - yc
struct __align__(16) Coef {
  float A, B, C, amp;
  float irf, fr, icf, fc;  // centre split into integer part and fraction
  int r0, r1, c0, c1;      // inclusive bounding box of the > tau footprint
};

struct Args {
  const double* params;
  const int32_t* zoom_pad;
  int n, slots, size;
  unsigned flags;
  void* out;
  int32_t* margin;
  double* peak;
  float* scratch;  // per-CTA raw-image slots when the image does not fit in smem
  int pitch;       // floats per stored raw row (size rounded up to 4)
  int stage_rows = 0;  // on-chip path: rows of each band that leave through a staged bulk copy
  // variable-size launch (on-chip path only): image i is sizes[i] x sizes[i] (all <= size), its
  // rows are stored (sizes[i] + 3) & ~3 elements apart starting at out + out_offs[i] elements
  // (offsets multiples of 4); raw_floats = floats of shared memory behind the tap table
  const int32_t* sizes = nullptr;
  const long long* out_offs = nullptr;
  int raw_floats = 0;
};

// ---- float64 evaluation of one raw pixel, operation order of gauss.py:77-88 ----
__device__ inline double eval64(const double* P, int slots, int r, int c) {
  double sum = 0.0;
  for (int k = 0; k < slots; ++k) {
    const double* p = P + k * RSB_JET_PARAM_STRIDE;
    const double amp = p[0];
    if (amp == 0.0) continue;
    const double cr = cos(p[5]), sr = sin(p[5]);
    const double ra = p[2] * cr - p[1] * sr;
    const double rb = p[2] * sr + p[1] * cr;
    const double ua = r * cr - c * sr;
    const double ub = r * sr + c * cr;
    const double a = (ra - ua) / p[4];
    const double b = (rb - ub) / p[3];
    sum += amp * exp(-(a * a + b * b) / 2.0);
  }
  return sum;
}

// ---- composite 4-tap operator of zoom_on_source followed by zoom_out -------------
// cv2.resize INTER_LINEAR per axis: f = (d+0.5)*S/D - 0.5, s = floor(f), w = f - s,
// clamped to [0, S-1] with w = 0 at the clamps.
struct Taps64 {
  int i[4];
  double w[4];
};
// sP = (double)(N + 2 pad) / (double)N and sC = (double)(N - 2 m) / (double)N are the same
// for every coordinate of an image: callers that build whole tables pass them in (two float64
// divisions less per entry; the FP64 pipe is narrow).
__device__ inline Taps64 composite_taps(int d, int N, int m, int pad, double sP, double sC);
__device__ inline Taps64 composite_taps(int d, int N, int m, int pad) {
  return composite_taps(d, N, m, pad, (double)(N + 2 * pad) / (double)N, (double)(N - 2 * m) / (double)N);
}
__device__ inline Taps64 composite_taps(int d, int N, int m, int pad, double sP, double sC) {
  Taps64 t;
  const int P = N + 2 * pad;
  const int Cs = N - 2 * m;
  double f = (d + 0.5) * sP - 0.5;
  int s = (int)floor(f);
  double w = f - (double)s;
  if (s < 0) { s = 0; w = 0.0; }
  if (s >= P - 1) { s = P - 1; w = 0.0; }
  const int s1 = min(s + 1, P - 1);
  const int z[2] = {s - pad, s1 - pad};
  const double wz[2] = {1.0 - w, w};
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    if (z[a] < 0 || z[a] >= N || wz[a] == 0.0) {
      t.i[2 * a] = t.i[2 * a + 1] = m;
      t.w[2 * a] = t.w[2 * a + 1] = 0.0;
      continue;
    }
    double g = (z[a] + 0.5) * sC - 0.5;
    int u = (int)floor(g);
    double v = g - (double)u;
    if (u < 0) { u = 0; v = 0.0; }
    if (u >= Cs - 1) { u = Cs - 1; v = 0.0; }
    const int u1 = min(u + 1, Cs - 1);
    t.i[2 * a] = m + u;
    t.i[2 * a + 1] = m + u1;
    t.w[2 * a] = wz[a] * (1.0 - v);
    t.w[2 * a + 1] = wz[a] * v;
  }
  return t;
}

// ---- warp 0: float64 parameter block -> compacted float32 coefficients ----------
template <class SM>
__device__ inline void build_coefs(SM& sm, const double* P, int slots, int N, int lane) {
  float amax = 0.f;
  for (int k = lane; k < slots; k += 32) amax = fmaxf(amax, fabsf((float)P[k * 6]));
  amax = warp_max(amax);
  const double tau = 1e-8 * (double)amax;  // image peak >= max amp: < 2e-7 of peak dropped
  int base = 0;
  for (int k0 = 0; k0 < slots; k0 += 32) {
    const int k = k0 + lane;
    bool live = false;
    Coef c;
    if (k < slots) {
      const double* p = P + k * 6;
      const double amp = p[0], x = p[1], y = p[2], sx = p[3], sy = p[4];
      if (amp != 0.0 && fabs(amp) > tau) {
        double sr, cr;
        sincos(p[5], &sr, &cr);
        // u = (dr cr - dc sr)/sy, v = (dr sr + dc cr)/sx ; exponent = -(u^2+v^2)/2 * log2(e)
        const double h = 0.5 * 1.4426950408889634;
        const double iu = 1.0 / (sy * sy), iv = 1.0 / (sx * sx);
        c.A = (float)(-h * (sr * sr * iu + cr * cr * iv));
        c.C = (float)(-h * (cr * cr * iu + sr * sr * iv));
        c.B = (float)(-2.0 * h * sr * cr * (iv - iu));
        c.amp = (float)amp;
        const double yr = rint(y), xr = rint(x);
        c.irf = (float)yr;
        c.fr = (float)(y - yr);
        c.icf = (float)xr;
        c.fc = (float)(x - xr);
        const double T = sqrt(2.0 * log(fabs(amp) / tau));
        const double er = T * sqrt(sy * sy * cr * cr + sx * sx * sr * sr);
        const double ec = T * sqrt(sy * sy * sr * sr + sx * sx * cr * cr);
        const double lo_r = fmax(floor(y - er), 0.0), hi_r = fmin(ceil(y + er), (double)(N - 1));
        const double lo_c = fmax(floor(x - ec), 0.0), hi_c = fmin(ceil(x + ec), (double)(N - 1));
        live = (lo_r <= hi_r) && (lo_c <= hi_c) && isfinite(er) && isfinite(ec);
        if (!(isfinite(er) && isfinite(ec))) {  // degenerate widths: keep, never cull
          live = true;
          c.r0 = 0; c.r1 = N - 1; c.c0 = 0; c.c1 = N - 1;
        } else if (live) {
          c.r0 = (int)lo_r; c.r1 = (int)hi_r; c.c0 = (int)lo_c; c.c1 = (int)hi_c;
        }
      }
    }
    const unsigned mask = __ballot_sync(0xffffffffu, live);
    if (live) sm.coef[base + __popc(mask & ((1u << lane) - 1u))] = c;
    base += __popc(mask);
  }
  if (lane == 0) {
    sm.nlive = base;
    sm.delta = 2e-6f * fmaxf(1.f, amax);
  }
}


// host-side launchers of the on-chip path (jet2.cu)
bool v2_supported(int size, int slots, int* cluster, size_t* smem_bytes);
int v2_launch(const Args& a, int cluster, size_t smem_bytes, cudaStream_t st);
// variable-size launch: a.size = largest size in the batch, a.sizes / a.out_offs per image
int v2_launch_var(const Args& a, cudaStream_t st);
// band renderer for 257 .. 512 pixels (jet3.cu): 8-CTA clusters, V pass through distributed shared memory
bool v3_supported(int size, int slots, size_t* smem_bytes);
int v3_launch(const Args& a, size_t smem_bytes, cudaStream_t st);

}  // namespace jet
}  // namespace rsb
