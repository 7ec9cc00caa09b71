// Fused jet renderer: splat -> margin -> zoom-in -> zoom-out -> max-normalise.
//
// One persistent CTA renders one image at a time.  Per image:
//   setup   : the [slots][6] float64 parameter block arrives by a 1-D bulk async
//             copy (TMA engine, double-buffered so the next image's block is in
//             flight while this one renders); warp 0 turns it into float32 conic
//             coefficients + bounding boxes (culling) for the live slots.
//   phase 1 : every warp renders (row, 128-column strip) items, 4 pixels per lane,
//             one MUFU.EX2 + 3 FFMA + 1 FADD per pixel-component, components whose
//             1e-8 footprint misses the strip are skipped warp-uniformly.  The raw
//             image is kept on chip (shared memory) or, when size*size*4 exceeds
//             shared memory, in an L2-resident per-CTA global scratch slot.
//   zoom    : (only when the host edge test says so) margin = min border distance
//             of pixels > 0.01 (borderline pixels re-evaluated in float64); the
//             two cv2.resize(INTER_LINEAR) calls of the reference collapse into ONE
//             separable 4-tap operator per axis, applied in place (H pass) and on
//             the fly (V pass, run twice: once for the maximum, once to write).
//   output  : every pixel is written exactly once with 128-bit stores, already
//             divided by the image maximum.
//
// Reference: src/radiosim/jets/jet.py:144-181, src/radiosim/utils/gauss.py:50-97,
// src/radiosim/utils/utils.py:90-177.
#include <cmath>
#include <mutex>
#include <vector>

#include "jet_common.cuh"

namespace rsb {
namespace jet {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kCtasPerSm = 2;

struct __align__(16) Fixed {
  double params[2][kMaxSlots * RSB_JET_PARAM_STRIDE];
  Coef coef[kMaxSlots];
  double red_d[kWarps];
  unsigned long long red_u[kWarps];
  uint64_t bar[2];
  double peak;
  double pix64[16];
  int red_i[kWarps];
  int nlive;
  int margin;
  unsigned long long best;
  float delta;
  float2 blk[kMaxSlots];  // per live slot {2^(2C), 2^(6C)} of the four-row blocks, or {-1, -1}: one exp per pixel
  float la[kMaxSlots];    // log2(amplitude) where the blocks are used
};

// 4 consecutive pixels of one row, all live components (culled per strip).
__device__ __forceinline__ void splat_row4(const Coef* __restrict__ coef, int nlive, int r,
                                           float rf, int strip_c0, float c0f, float acc[4]) {
  const int strip_c1 = strip_c0 + kStrip - 1;
  for (int k = 0; k < nlive; ++k) {
    const Coef c = coef[k];
    if (r < c.r0 || r > c.r1 || strip_c1 < c.c0 || strip_c0 > c.c1) continue;  // warp-uniform
    const float dr = (rf - c.irf) - c.fr;
    const float dc = (c0f - c.icf) - c.fc;
    const float Bq = c.B * dr;
    const float Cq = c.C * dr * dr;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float d = dc + (float)j;
      const float t = fmaf(fmaf(c.A, d, Bq), d, Cq);
      acc[j] = fmaf(c.amp, ex2(t), acc[j]);
    }
  }
}

// One live slot over an 8-row x 64-column tile, two adjacent pixels per lane (packed FP32x2).  Rows in
// blocks of four from two exponentials per column,
//   E(i) = E0 U^i 2^(C (i^2 - i)),  E0 = g(r_b, c),  U = 2^(B dc + 2 C dr_b + C)
// (half the MUFU work of one exponential per pixel; same scheme as jet2.cu), unless the slot does not
// suit it (amplitude <= 0, sigma below ~0.6 px, non-finite): then one exponential per pixel.
__device__ __forceinline__ void tile_slot(f2 acc2[8], const Coef& c, float la, float2 kk, float r0f, float c0f) {
  const float d0 = (c0f - c.icf) - c.fc;
  const f2 d = f2_pack(d0, d0 + 1.f);
  const f2 Q = f2_mul(f2_pack(c.B, c.B), d);
  const float dr0 = (r0f - c.irf) - c.fr;
  if (kk.x > 0.f) {
    const f2 Pq = f2_fma(f2_mul(f2_pack(c.A, c.A), d), d, f2_pack(la, la));  // A dc^2 + log2(amp)
    const f2 k2 = f2_pack(kk.x, kk.x), k3 = f2_pack(kk.y, kk.y);
#pragma unroll
    for (int b = 0; b < 8; b += 4) {
      const float dr = dr0 + (float)b;
      const float Cq = c.C * dr * dr;
      const float tw = fmaf(2.f * c.C, dr, c.C);  // 2 C dr + C
      float ta, tb, wa, wb;
      f2_unpack(f2_fma(f2_pack(dr, dr), Q, f2_add(Pq, f2_pack(Cq, Cq))), ta, tb);
      f2_unpack(f2_add(Q, f2_pack(tw, tw)), wa, wb);
      // (the clamp never binds where the block holds a pixel above the cut-off; elsewhere it keeps U^3 finite)
      const f2 E0 = f2_pack(ex2(ta), ex2(tb));
      const f2 U = f2_pack(ex2(fminf(wa, 40.f)), ex2(fminf(wb, 40.f)));
      const f2 U2 = f2_mul(U, U);
      acc2[b] = f2_add(acc2[b], E0);
      acc2[b + 1] = f2_fma(E0, U, acc2[b + 1]);
      acc2[b + 2] = f2_fma(E0, f2_mul(U2, k2), acc2[b + 2]);
      acc2[b + 3] = f2_fma(E0, f2_mul(f2_mul(U2, U), k3), acc2[b + 3]);
    }
  } else {
    const f2 Pq = f2_mul(f2_mul(f2_pack(c.A, c.A), d), d);
    const f2 amp2 = f2_pack(c.amp, c.amp);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float dr = dr0 + (float)i;
      const float Cq = c.C * dr * dr;
      float ta, tb;
      f2_unpack(f2_fma(f2_pack(dr, dr), Q, f2_add(Pq, f2_pack(Cq, Cq))), ta, tb);
      acc2[i] = f2_fma(amp2, f2_pack(ex2(ta), ex2(tb)), acc2[i]);
    }
  }
}

template <typename OutT>
__device__ __forceinline__ void store4(OutT* row, int c0, int N, bool vec, float4 v) {
  if (vec) {
    if constexpr (sizeof(OutT) == 4) {
      *reinterpret_cast<float4*>(row + c0) = v;
    } else {
      double2* p = reinterpret_cast<double2*>(row + c0);
      p[0] = make_double2((double)v.x, (double)v.y);
      p[1] = make_double2((double)v.z, (double)v.w);
    }
  } else {
    const float a[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (c0 + j < N) row[c0 + j] = (OutT)a[j];
  }
}

template <bool kSmemRaw, int NS, typename OutT>
__global__ void __launch_bounds__(kThreads, kCtasPerSm) jet_kernel(const Args a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Fixed& sm = *reinterpret_cast<Fixed*>(smem_raw);
  const int N = a.size, pitch = a.pitch, slots = a.slots;
  const int npad = (N + 3) & ~3;
  float* tab_w = reinterpret_cast<float*>(smem_raw + sizeof(Fixed));  // [4][npad]
  int* tab_i = reinterpret_cast<int*>(tab_w + 4 * npad);              // [4][npad]
  float* raw;
  if constexpr (kSmemRaw)
    raw = reinterpret_cast<float*>(tab_i + 4 * npad);
  else
    raw = a.scratch + (size_t)blockIdx.x * (size_t)pitch * (size_t)N;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t pbytes = (uint32_t)slots * RSB_JET_PARAM_STRIDE * 8u;
  const int strips = (N + kStrip - 1) / kStrip;
  const int items = N * strips;
  const bool vec = (N & 3) == 0 && (reinterpret_cast<uintptr_t>(a.out) & 15) == 0;

  if (tid == 0) {
    mbar_init(&sm.bar[0], 1);
    mbar_init(&sm.bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0 && (int)blockIdx.x < a.n) {
    mbar_expect_tx(&sm.bar[0], pbytes);
    bulk_g2s(sm.params[0], a.params + (size_t)blockIdx.x * slots * 6, pbytes, &sm.bar[0]);
  }

  int it = 0;
  for (int img = blockIdx.x; img < a.n; img += gridDim.x, ++it) {
    const int buf = it & 1;
    // prefetch the next image's parameter block into the other buffer
    if (tid == 0 && img + (int)gridDim.x < a.n) {
      fence_proxy_async();
      mbar_expect_tx(&sm.bar[buf ^ 1], pbytes);
      bulk_g2s(sm.params[buf ^ 1], a.params + (size_t)(img + gridDim.x) * slots * 6, pbytes,
               &sm.bar[buf ^ 1]);
    }
    mbar_wait(&sm.bar[buf], (it >> 1) & 1);
    const double* P = sm.params[buf];
    const int zpad = a.zoom_pad[img];
    const bool zoom = zpad >= 0;
    const bool normalize = (a.flags & RSB_JET_NORMALIZE) != 0;

    if (warp == 0) {
      build_coefs(sm, P, slots, N, lane);
      __syncwarp();
      bool nonneg = true;
      float amax = 0.f;
      for (int k = lane; k < slots; k += 32) {
        nonneg = nonneg && !(P[k * 6] < 0.0);
        amax = fmaxf(amax, fabsf((float)P[k * 6]));
      }
      nonneg = __all_sync(0xffffffffu, nonneg);
      amax = warp_max(amax);
      for (int k = lane; k < sm.nlive; k += 32) {
        const Coef c = sm.coef[k];
        const bool blocks = c.amp > 0.f && c.C >= -2.f && isfinite(c.A) && isfinite(c.B) && isfinite(c.C);
        sm.la[k] = blocks ? log2f(c.amp) : 0.f;
        sm.blk[k] = blocks ? make_float2(exp2f(2.f * c.C), exp2f(6.f * c.C)) : make_float2(-1.f, -1.f);
      }
      // float32 noise of a rendered pixel near the 0.01 threshold (see jet2.cu: relative to the pixel value
      // when every amplitude is >= 0, plus the tails the footprint cut drops; else relative to the largest one)
      if (lane == 0) sm.delta = nonneg ? 3e-7f + 4e-7f * amax : 6e-6f * fmaxf(1.f, amax);
    }
    __syncthreads();
    const int nlive = sm.nlive;
    const float thr = (float)kEdgeThr, delta = sm.delta;

    // ---------------- phase 1: splat ------------------------------------------
    int my_margin = N;                  // zoom branch: min border distance of px > thr
    unsigned long long my_best = 0ull;  // else: (value bits << 32) | ~pixel index
    // 8 x 64 tiles, two adjacent pixels per lane; slots culled per tile by their > tau bounding box
    const int tcols = (N + 63) >> 6, ntiles = ((N + 7) >> 3) * tcols;
    for (int t = warp; t < ntiles; t += kWarps) {
      const int g = t / tcols, tc0 = (t - g * tcols) << 6;
      const int r0 = g << 3, c0 = tc0 + 2 * lane, rend = min(N, r0 + 8);
      f2 acc2[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc2[i] = 0ull;
      for (int k = 0; k < nlive; ++k) {
        const Coef c = sm.coef[k];
        if (rend - 1 < c.r0 || r0 > c.r1 || tc0 + 63 < c.c0 || tc0 > c.c1) continue;  // warp-uniform
        tile_slot(acc2, c, sm.la[k], sm.blk[k], (float)r0, (float)c0);
      }
      if (c0 < N) {
        const bool second = c0 + 1 < N;   // odd sizes: the pair's second column lies outside the image
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = r0 + i;
          if (r >= rend) break;
          float v[2];
          f2_unpack(acc2[i], v[0], v[1]);
          if (!second) v[1] = 0.f;
          *reinterpret_cast<float2*>(raw + (size_t)r * pitch + c0) = make_float2(v[0], v[1]);
          if (zoom) {
            if (fmaxf(v[0], v[1]) > thr - delta) {
              const int dr = min(r, N - 1 - r);
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                const int c = c0 + j;
                const int d = min(dr, min(c, N - 1 - c));
                if (c < N && d < my_margin && v[j] > thr - delta) {
                  bool above = v[j] > thr;
                  if (v[j] < thr + delta) above = eval64(P, slots, r, c) > kEdgeThr;
                  if (above) my_margin = d;
                }
              }
            }
          } else if (normalize) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const int c = c0 + j;
              if (c < N) {
                const unsigned long long key =
                    ((unsigned long long)__float_as_uint(fmaxf(v[j], 0.f)) << 32) |
                    (unsigned)(~(unsigned)(r * N + c));
                my_best = key > my_best ? key : my_best;
              }
            }
          }
        }
      }
    }

    double peak = 1.0;
    int margin = 0;
    OutT* out = reinterpret_cast<OutT*>(a.out) + (size_t)img * N * N;

    if (!zoom) {
      // ------------- plain branch: (normalise and) write the raw image ----------
      if (normalize) {
        my_best = warp_max_u64(my_best);
        if (lane == 0) sm.red_u[warp] = my_best;
        __syncthreads();
        if (warp == 0) {
          unsigned long long b = lane < kWarps ? sm.red_u[lane] : 0ull;
          b = warp_max_u64(b);
          if (lane == 0) {
            const unsigned idx = ~(unsigned)(b & 0xffffffffull);
            sm.peak = eval64(P, slots, (int)(idx / (unsigned)N), (int)(idx % (unsigned)N));
          }
        }
        __syncthreads();
        peak = sm.peak;
      } else {
        __syncthreads();
      }
      const float inv = normalize ? (float)(1.0 / peak) : 1.f;
      const int q = npad >> 2;
      for (int e = tid; e < N * q; e += kThreads) {
        const int r = e / q, c0 = (e - r * q) << 2;
        float4 v = *reinterpret_cast<const float4*>(raw + (size_t)r * pitch + c0);
        v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
        store4<OutT>(out + (size_t)r * N, c0, N, vec, v);
      }
    } else {
      // ------------- zoom branch ------------------------------------------------
      my_margin = warp_min(my_margin);
      if (lane == 0) sm.red_i[warp] = my_margin;
      __syncthreads();
      if (warp == 0) {
        int m = lane < kWarps ? sm.red_i[lane] : N;
        m = warp_min(m);
        if (lane == 0) sm.margin = (m >= N) ? 0 : m;  // empty mask -> idx 0 (argmax of zeros)
      }
      __syncthreads();
      margin = sm.margin;
      for (int d = tid; d < npad; d += kThreads) {
        if (d < N) {
          const Taps64 t = composite_taps(d, N, margin, zpad);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            tab_w[k * npad + d] = (float)t.w[k];
            tab_i[k * npad + d] = t.i[k];
          }
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            tab_w[k * npad + d] = 0.f;
            tab_i[k * npad + d] = margin;
          }
        }
      }
      __syncthreads();

      // H pass, in place: row r' of raw becomes row r' of T (all N columns).
      for (int r = margin + warp; r < N - margin; r += kWarps) {
        float* row = raw + (size_t)r * pitch;
        float4 res[NS];
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          const int c0 = s * kStrip + 4 * lane;
          res[s] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (c0 < N) {
            float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float4 w = *reinterpret_cast<const float4*>(tab_w + k * npad + c0);
              const int4 i = *reinterpret_cast<const int4*>(tab_i + k * npad + c0);
              o[0] = fmaf(w.x, row[i.x], o[0]);
              o[1] = fmaf(w.y, row[i.y], o[1]);
              o[2] = fmaf(w.z, row[i.z], o[2]);
              o[3] = fmaf(w.w, row[i.w], o[3]);
            }
            res[s] = make_float4(o[0], o[1], o[2], o[3]);
          }
        }
        __syncwarp();
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          const int c0 = s * kStrip + 4 * lane;
          if (c0 < N) *reinterpret_cast<float4*>(row + c0) = res[s];
        }
      }
      __syncthreads();

      // V pass A: maximum (and where) of the resampled image.
      float inv = 1.f;
      if (normalize) {
        my_best = 0ull;
        for (int item = warp; item < items; item += kWarps) {
          const int r = item / strips, s = item - r * strips;
          const int c0 = s * kStrip + 4 * lane;
          const float w0 = tab_w[r], w1 = tab_w[npad + r], w2 = tab_w[2 * npad + r],
                      w3 = tab_w[3 * npad + r];
          if (c0 >= N || (w0 == 0.f && w1 == 0.f && w2 == 0.f && w3 == 0.f)) continue;
          const float4 t0 = *reinterpret_cast<const float4*>(raw + (size_t)tab_i[r] * pitch + c0);
          const float4 t1 = *reinterpret_cast<const float4*>(raw + (size_t)tab_i[npad + r] * pitch + c0);
          const float4 t2 = *reinterpret_cast<const float4*>(raw + (size_t)tab_i[2 * npad + r] * pitch + c0);
          const float4 t3 = *reinterpret_cast<const float4*>(raw + (size_t)tab_i[3 * npad + r] * pitch + c0);
          const float v[4] = {fmaf(w3, t3.x, fmaf(w2, t2.x, fmaf(w1, t1.x, w0 * t0.x))),
                              fmaf(w3, t3.y, fmaf(w2, t2.y, fmaf(w1, t1.y, w0 * t0.y))),
                              fmaf(w3, t3.z, fmaf(w2, t2.z, fmaf(w1, t1.z, w0 * t0.z))),
                              fmaf(w3, t3.w, fmaf(w2, t2.w, fmaf(w1, t1.w, w0 * t0.w)))};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = c0 + j;
            if (c < N) {
              const unsigned long long key =
                  ((unsigned long long)__float_as_uint(fmaxf(v[j], 0.f)) << 32) |
                  (unsigned)(~(unsigned)(r * N + c));
              my_best = key > my_best ? key : my_best;
            }
          }
        }
        my_best = warp_max_u64(my_best);
        if (lane == 0) sm.red_u[warp] = my_best;
        __syncthreads();
        if (warp == 0) {
          unsigned long long b = lane < kWarps ? sm.red_u[lane] : 0ull;
          b = warp_max_u64(b);
          // float64 value of the arg-max pixel: 4x4 raw pixels, float64 weights
          const unsigned idx = ~(unsigned)(b & 0xffffffffull);
          const int pr = (int)(idx / (unsigned)N), pc = (int)(idx % (unsigned)N);
          const Taps64 tr = composite_taps(pr, N, margin, zpad);
          const Taps64 tc = composite_taps(pc, N, margin, zpad);
          if (lane < 16) {
            const int i = lane >> 2, j = lane & 3;
            const double w = tr.w[i] * tc.w[j];
            sm.pix64[lane] = (w != 0.0) ? eval64(P, slots, tr.i[i], tc.i[j]) : 0.0;
          }
          __syncwarp();
          if (lane == 0) {
            double acc = 0.0;
            for (int i = 0; i < 4; ++i) {
              double h = 0.0;
              for (int j = 0; j < 4; ++j) h += tc.w[j] * sm.pix64[i * 4 + j];
              acc += tr.w[i] * h;
            }
            sm.peak = acc;
          }
        }
        __syncthreads();
        peak = sm.peak;
        inv = (float)(1.0 / peak);
      }

      // V pass B: recompute, scale, write every output pixel once.
      for (int item = warp; item < items; item += kWarps) {
        const int r = item / strips, s = item - r * strips;
        const int c0 = s * kStrip + 4 * lane;
        if (c0 >= N) continue;
        const float w0 = tab_w[r] * inv, w1 = tab_w[npad + r] * inv, w2 = tab_w[2 * npad + r] * inv,
                    w3 = tab_w[3 * npad + r] * inv;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (!(w0 == 0.f && w1 == 0.f && w2 == 0.f && w3 == 0.f)) {
          const float4 t0 = *reinterpret_cast<const float4*>(raw + (size_t)tab_i[r] * pitch + c0);
          const float4 t1 = *reinterpret_cast<const float4*>(raw + (size_t)tab_i[npad + r] * pitch + c0);
          const float4 t2 = *reinterpret_cast<const float4*>(raw + (size_t)tab_i[2 * npad + r] * pitch + c0);
          const float4 t3 = *reinterpret_cast<const float4*>(raw + (size_t)tab_i[3 * npad + r] * pitch + c0);
          v.x = fmaf(w3, t3.x, fmaf(w2, t2.x, fmaf(w1, t1.x, w0 * t0.x)));
          v.y = fmaf(w3, t3.y, fmaf(w2, t2.y, fmaf(w1, t1.y, w0 * t0.y)));
          v.z = fmaf(w3, t3.z, fmaf(w2, t2.z, fmaf(w1, t1.z, w0 * t0.z)));
          v.w = fmaf(w3, t3.w, fmaf(w2, t2.w, fmaf(w1, t1.w, w0 * t0.w)));
        }
        store4<OutT>(out + (size_t)r * N, c0, N, vec, v);
      }
    }
    if (tid == 0) {
      a.margin[img] = margin;
      a.peak[img] = peak;
    }
    __syncthreads();  // raw / tab / coef / params[buf] are reused by the next image
  }
}

// ------------------------------------------------------------------ host side
struct Plan {
  bool smem_raw;
  int ns;
  size_t smem_bytes;
  int grid_cap;  // CTAs that can be resident
  int pitch;
};

static size_t fixed_bytes(int size) {
  const int npad = (size + 3) & ~3;
  return sizeof(Fixed) + (size_t)npad * 32;
}

static int make_plan(int size, Plan* p) {
  int dev = 0, max_optin = 0;
  RSB_CUDA(cudaGetDevice(&dev));
  RSB_CUDA(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  p->pitch = (size + 3) & ~3;
  const size_t raw_bytes = (size_t)p->pitch * size * 4;
  const size_t fx = fixed_bytes(size);
  p->smem_raw = fx + raw_bytes <= (size_t)max_optin;
  p->smem_bytes = fx + (p->smem_raw ? raw_bytes : 0);
  const int strips = (size + kStrip - 1) / kStrip;
  p->ns = strips <= 1 ? 1 : strips <= 2 ? 2 : strips <= 4 ? 4 : 8;
  // per-SM residency: threads (2048/512 = 4, capped by launch bounds) and shared memory
  int per_sm = kCtasPerSm;
  const size_t sm_total = 228 * 1024;
  const int by_smem = (int)(sm_total / (p->smem_bytes + 1024));
  if (by_smem < per_sm) per_sm = by_smem < 1 ? 1 : by_smem;
  p->grid_cap = sm_count() * per_sm;
  return 0;
}

template <bool S, int NS, typename T>
static int launch_one(const Args& a, const Plan& p, int grid, cudaStream_t st) {
  auto k = jet_kernel<S, NS, T>;
  RSB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem_bytes));
  k<<<grid, kThreads, p.smem_bytes, st>>>(a);
  RSB_LAUNCHED();
  return 0;
}

template <bool S, typename T>
static int launch_ns(const Args& a, const Plan& p, int grid, cudaStream_t st) {
  switch (p.ns) {
    case 1: return launch_one<S, 1, T>(a, p, grid, st);
    case 2: return launch_one<S, 2, T>(a, p, grid, st);
    case 4: return launch_one<S, 4, T>(a, p, grid, st);
    default: return launch_one<S, 8, T>(a, p, grid, st);
  }
}

static int launch(const Args& a, const Plan& p, int grid, cudaStream_t st) {
  const bool f64 = (a.flags & RSB_JET_OUT_F64) != 0;
  if (p.smem_raw) return f64 ? launch_ns<true, double>(a, p, grid, st) : launch_ns<true, float>(a, p, grid, st);
  return f64 ? launch_ns<false, double>(a, p, grid, st) : launch_ns<false, float>(a, p, grid, st);
}

// ---------------------------------------------------------------- plane kernel
// training_type == "gauss": slot planes after zoom_on_source, divided by the peak.
// Each output pixel is the bilinear blend of <= 4 analytic Gaussian samples.
struct PlaneArgs {
  const double* params;
  const int32_t* zoom_pad;
  const int32_t* margin;
  const double* peak;
  int n, slots, size;
  unsigned flags;
  void* out;
};

template <typename OutT>
__global__ void __launch_bounds__(256) planes_kernel(const PlaneArgs a) {
  // grid: (row blocks, slot, image)
  const int img = blockIdx.z, k = blockIdx.y, N = a.size;
  const double* p = a.params + ((size_t)img * a.slots + k) * 6;
  OutT* out = reinterpret_cast<OutT*>(a.out) + ((size_t)img * a.slots + k) * N * N;
  const double amp = p[0];
  const bool zoom = a.zoom_pad[img] >= 0;
  const int m = a.margin[img];
  const float inv = (a.flags & RSB_JET_NORMALIZE) ? (float)(1.0 / a.peak[img]) : 1.f;
  float A = 0.f, B = 0.f, C = 0.f, xc = 0.f, yc = 0.f, fxc = 0.f, fyc = 0.f;
  if (amp != 0.0) {
    double sr, cr;
    sincos(p[5], &sr, &cr);
    const double h = 0.5 * 1.4426950408889634;
    const double iu = 1.0 / (p[4] * p[4]), iv = 1.0 / (p[3] * p[3]);
    A = (float)(-h * (sr * sr * iu + cr * cr * iv));
    C = (float)(-h * (cr * cr * iu + sr * sr * iv));
    B = (float)(-2.0 * h * sr * cr * (iv - iu));
    xc = (float)rint(p[1]); fxc = (float)(p[1] - rint(p[1]));
    yc = (float)rint(p[2]); fyc = (float)(p[2] - rint(p[2]));
  }
  const float ampf = (float)amp * inv;
  const int Cs = N - 2 * m;
  const double sc = (double)Cs / (double)N;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < N * N; e += gridDim.x * blockDim.x) {
    const int r = e / N, c = e - r * N;
    float v = 0.f;
    if (amp != 0.0) {
      int ri[2] = {r, r}, ci[2] = {c, c};
      float wr = 0.f, wc = 0.f;
      if (zoom) {
        double g = (r + 0.5) * sc - 0.5;
        int u = (int)floor(g);
        double w = g - u;
        if (u < 0) { u = 0; w = 0; }
        if (u >= Cs - 1) { u = Cs - 1; w = 0; }
        ri[0] = m + u; ri[1] = m + min(u + 1, Cs - 1); wr = (float)w;
        g = (c + 0.5) * sc - 0.5;
        u = (int)floor(g);
        w = g - u;
        if (u < 0) { u = 0; w = 0; }
        if (u >= Cs - 1) { u = Cs - 1; w = 0; }
        ci[0] = m + u; ci[1] = m + min(u + 1, Cs - 1); wc = (float)w;
      }
      float s[2][2];
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const float dr = ((float)ri[i] - yc) - fyc, dc = ((float)ci[j] - xc) - fxc;
          s[i][j] = ex2(fmaf(fmaf(A, dc, B * dr), dc, C * dr * dr));
        }
      const float top = fmaf(wc, s[0][1] - s[0][0], s[0][0]);
      const float bot = fmaf(wc, s[1][1] - s[1][0], s[1][0]);
      v = ampf * fmaf(wr, bot - top, top);
    }
    out[e] = (OutT)v;
  }
}

}  // namespace jet
}  // namespace rsb

using namespace rsb;
using namespace rsb::jet;

static int check_common(const char* fn, int n, int slots, int size) {
  RSB_ARG(n >= 0, "%s: n must be >= 0 (got %d)", fn, n);
  RSB_ARG(slots >= 1 && slots <= kMaxSlots, "%s: slots must be in [1, %d] (got %d)", fn, kMaxSlots, slots);
  RSB_ARG(size >= 2 && size <= 1024, "%s: size must be in [2, 1024] (got %d)", fn, size);
  return 0;
}

// sizes 257 .. 512 on the band renderer of jet3.cu (RSB_NO_BANDS=1: the scratch-image kernel below, for comparison)
static bool bands_supported(int size, int slots, size_t* smem) {
  static const bool off = [] { const char* e = getenv("RSB_NO_BANDS"); return e && e[0] == '1'; }();
  return !off && v3_supported(size, slots, smem);
}
static int bands_launch(const double* params, const int32_t* zoom_pad, int n, int slots, int size, unsigned flags,
                        void* out_img, int32_t* out_margin, double* out_peak, size_t smem, void* stream) {
  Args a3;
  a3.params = params; a3.zoom_pad = zoom_pad; a3.n = n; a3.slots = slots; a3.size = size;
  a3.flags = flags; a3.out = out_img; a3.margin = out_margin; a3.peak = out_peak;
  a3.scratch = nullptr; a3.pitch = (size + 3) & ~3;
  return v3_launch(a3, smem, (cudaStream_t)stream);
}

extern "C" {

int rsb_jet_workspace_bytes(int n, int slots, int size, size_t* bytes) {
  RSB_ARG(bytes != nullptr, "rsb_jet_workspace_bytes: bytes is NULL");
  if (int e = check_common("rsb_jet_workspace_bytes", n, slots, size)) return e;
  {
    int cluster = 1;
    size_t smem = 0;
    if (v2_supported(size, slots, &cluster, &smem) || bands_supported(size, slots, &smem)) {
      *bytes = 0;
      return 0;
    }
  }
  Plan p;
  if (int e = make_plan(size, &p)) return e;
  if (p.smem_raw) {
    *bytes = 0;
  } else {
    const int grid = n < p.grid_cap ? (n > 0 ? n : 1) : p.grid_cap;
    *bytes = (size_t)grid * p.pitch * size * sizeof(float);
  }
  return 0;
}

int rsb_jet_render(const double* params, const int32_t* zoom_pad, int n, int slots, int size,
                   unsigned flags, void* out_img, int32_t* out_margin, double* out_peak,
                   void* workspace, size_t workspace_bytes, void* stream) {
  if (int e = check_common("rsb_jet_render", n, slots, size)) return e;
  if (n == 0) return 0;
  RSB_ARG(params && zoom_pad && out_img && out_margin && out_peak, "rsb_jet_render: NULL buffer");
  RSB_ARG(((uintptr_t)params & 15) == 0, "rsb_jet_render: params must be 16-byte aligned");
  {
    int cluster = 1;
    size_t smem = 0;
    if (v2_supported(size, slots, &cluster, &smem)) {  // on-chip path: no scratch at all
      Args a2;
      a2.params = params; a2.zoom_pad = zoom_pad; a2.n = n; a2.slots = slots; a2.size = size;
      a2.flags = flags; a2.out = out_img; a2.margin = out_margin; a2.peak = out_peak;
      a2.scratch = nullptr; a2.pitch = (size + 3) & ~3;
      return v2_launch(a2, cluster, smem, (cudaStream_t)stream);
    }
    if (bands_supported(size, slots, &smem))
      return bands_launch(params, zoom_pad, n, slots, size, flags, out_img, out_margin, out_peak, smem, stream);
  }
  Plan p;
  if (int e = make_plan(size, &p)) return e;
  const int grid = n < p.grid_cap ? n : p.grid_cap;
  if (!p.smem_raw) {
    const size_t need = (size_t)grid * p.pitch * size * sizeof(float);
    RSB_ARG(workspace != nullptr && workspace_bytes >= need,
            "rsb_jet_render: workspace too small (%zu < %zu bytes)", workspace_bytes, need);
  }
  Args a;
  a.params = params; a.zoom_pad = zoom_pad; a.n = n; a.slots = slots; a.size = size;
  a.flags = flags; a.out = out_img; a.margin = out_margin; a.peak = out_peak;
  a.scratch = (float*)workspace; a.pitch = p.pitch;
  return launch(a, p, grid, (cudaStream_t)stream);
}

int rsb_jet_render_var(const double* params, const int32_t* zoom_pad, const int32_t* sizes,
                       const long long* out_offs, int n, int slots, int max_size, unsigned flags,
                       void* out_img, int32_t* out_margin, double* out_peak, void* stream) {
  if (int e = check_common("rsb_jet_render_var", n, slots, max_size)) return e;
  if (n == 0) return 0;
  RSB_ARG(params && zoom_pad && sizes && out_offs && out_img && out_margin && out_peak,
          "rsb_jet_render_var: NULL buffer");
  RSB_ARG(((uintptr_t)params & 15) == 0 && ((uintptr_t)out_img & 15) == 0,
          "rsb_jet_render_var: params and out_img must be 16-byte aligned");
  RSB_ARG(max_size <= 256 && slots <= 32,
          "rsb_jet_render_var: sizes up to 256 and up to 32 slots (got %d, %d)", max_size, slots);
  Args a;
  a.params = params; a.zoom_pad = zoom_pad; a.n = n; a.slots = slots; a.size = max_size;
  a.flags = flags; a.out = out_img; a.margin = out_margin; a.peak = out_peak;
  a.scratch = nullptr; a.pitch = (max_size + 3) & ~3;
  a.sizes = sizes; a.out_offs = out_offs;
  return v2_launch_var(a, (cudaStream_t)stream);
}

int rsb_jet_planes(const double* params, const int32_t* zoom_pad, const int32_t* margin,
                   const double* peak, int n, int slots, int size, unsigned flags, void* out,
                   void* stream) {
  if (int e = check_common("rsb_jet_planes", n, slots, size)) return e;
  if (n == 0) return 0;
  RSB_ARG(params && zoom_pad && margin && peak && out, "rsb_jet_planes: NULL buffer");
  PlaneArgs a{params, zoom_pad, margin, peak, n, slots, size, flags, out};
  int bx = (size * size + 255) / 256;
  if (bx > 64) bx = 64;
  // grid.z is limited to 65535 images per launch
  for (int i0 = 0; i0 < n; i0 += 65535) {
    const int cnt = (n - i0) < 65535 ? (n - i0) : 65535;
    PlaneArgs b = a;
    b.params += (size_t)i0 * slots * 6;
    b.zoom_pad += i0; b.margin += i0; b.peak += i0;
    const size_t esz = (flags & RSB_JET_OUT_F64) ? 8 : 4;
    b.out = (char*)out + (size_t)i0 * slots * size * size * esz;
    dim3 grid(bx, slots, cnt);
    if (flags & RSB_JET_OUT_F64)
      planes_kernel<double><<<grid, 256, 0, (cudaStream_t)stream>>>(b);
    else
      planes_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(b);
    RSB_LAUNCHED();
  }
  return 0;
}


// ---- host-buffer convenience: chunked, 3 streams, copy/compute overlap ------------
namespace {
struct HostCtx {
  static constexpr int kLanes = 3;
  cudaStream_t st[kLanes] = {nullptr, nullptr, nullptr};
  void* params[kLanes] = {nullptr, nullptr, nullptr};
  void* pad[kLanes] = {nullptr, nullptr, nullptr};
  void* out[kLanes] = {nullptr, nullptr, nullptr};
  void* margin[kLanes] = {nullptr, nullptr, nullptr};
  void* peak[kLanes] = {nullptr, nullptr, nullptr};
  void* work[kLanes] = {nullptr, nullptr, nullptr};
  size_t cap_params = 0, cap_out = 0, cap_n = 0, cap_work = 0;
};
std::mutex g_host_mu[64];  // one per device: host threads driving different GPUs do not wait on each other
HostCtx g_host_ctx[64];

int grow(void** p, size_t bytes) {
  if (*p) RSB_CUDA(cudaFree(*p));
  *p = nullptr;
  RSB_CUDA(cudaMalloc(p, bytes ? bytes : 16));
  // test aid (same switch as the Python block pool): NaN bit patterns for whoever reads before writing
  static const bool poison = [] { const char* e = getenv("RSB_POISON_DEVICE_BLOCKS"); return e && e[0] == '1'; }();
  if (poison) RSB_CUDA(cudaMemset(*p, 0xFF, bytes ? bytes : 16));
  return 0;
}
}  // namespace

int rsb_jet_render_host(const double* params, const int32_t* zoom_pad, int n, int slots, int size,
                        unsigned flags, void* out_img, int32_t* out_margin, double* out_peak) {
  if (int e = check_common("rsb_jet_render_host", n, slots, size)) return e;
  if (n == 0) return 0;
  RSB_ARG(params && zoom_pad && out_img && out_margin && out_peak, "rsb_jet_render_host: NULL buffer");
  int dev = 0;
  RSB_CUDA(cudaGetDevice(&dev));
  RSB_ARG(dev >= 0 && dev < 64, "rsb_jet_render_host: device index %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_host_mu[dev]);
  HostCtx& cx = g_host_ctx[dev];
  const size_t esz = (flags & RSB_JET_OUT_F64) ? 8 : 4;
  const size_t img_bytes = (size_t)size * size * esz;
  const size_t par_bytes = (size_t)slots * 6 * sizeof(double);
  size_t chunk = ((size_t)96 << 20) / img_bytes;
  if (chunk < 320) chunk = 320;
  if (chunk > 16384) chunk = 16384;
  if (chunk > (size_t)n) chunk = (size_t)n;
  size_t work_bytes = 0;
  if (int e = rsb_jet_workspace_bytes((int)chunk, slots, size, &work_bytes)) return e;
  for (int l = 0; l < HostCtx::kLanes; ++l)
    if (!cx.st[l]) RSB_CUDA(cudaStreamCreateWithFlags(&cx.st[l], cudaStreamNonBlocking));
  if (cx.cap_params < chunk * par_bytes) {
    for (int l = 0; l < HostCtx::kLanes; ++l)
      if (int e = grow(&cx.params[l], chunk * par_bytes)) return e;
    cx.cap_params = chunk * par_bytes;
  }
  if (cx.cap_out < chunk * img_bytes) {
    for (int l = 0; l < HostCtx::kLanes; ++l)
      if (int e = grow(&cx.out[l], chunk * img_bytes)) return e;
    cx.cap_out = chunk * img_bytes;
  }
  if (cx.cap_n < chunk) {
    for (int l = 0; l < HostCtx::kLanes; ++l) {
      if (int e = grow(&cx.pad[l], chunk * 4)) return e;
      if (int e = grow(&cx.margin[l], chunk * 4)) return e;
      if (int e = grow(&cx.peak[l], chunk * 8)) return e;
    }
    cx.cap_n = chunk;
  }
  if (cx.cap_work < work_bytes) {
    for (int l = 0; l < HostCtx::kLanes; ++l)
      if (int e = grow(&cx.work[l], work_bytes)) return e;
    cx.cap_work = work_bytes;
  }
  int lane = 0;
  for (size_t i0 = 0; i0 < (size_t)n; i0 += chunk, lane = (lane + 1) % HostCtx::kLanes) {
    const size_t cnt = ((size_t)n - i0) < chunk ? ((size_t)n - i0) : chunk;
    cudaStream_t st = cx.st[lane];
    RSB_CUDA(cudaMemcpyAsync(cx.params[lane], params + i0 * slots * 6, cnt * par_bytes,
                             cudaMemcpyHostToDevice, st));
    RSB_CUDA(cudaMemcpyAsync(cx.pad[lane], zoom_pad + i0, cnt * 4, cudaMemcpyHostToDevice, st));
    if (int e = rsb_jet_render((const double*)cx.params[lane], (const int32_t*)cx.pad[lane], (int)cnt,
                               slots, size, flags, cx.out[lane], (int32_t*)cx.margin[lane],
                               (double*)cx.peak[lane], cx.work[lane], cx.cap_work, (void*)st))
      return e;
    RSB_CUDA(cudaMemcpyAsync((char*)out_img + i0 * img_bytes, cx.out[lane], cnt * img_bytes,
                             cudaMemcpyDeviceToHost, st));
    RSB_CUDA(cudaMemcpyAsync(out_margin + i0, cx.margin[lane], cnt * 4, cudaMemcpyDeviceToHost, st));
    RSB_CUDA(cudaMemcpyAsync(out_peak + i0, cx.peak[lane], cnt * 8, cudaMemcpyDeviceToHost, st));
  }
  for (int l = 0; l < HostCtx::kLanes; ++l) RSB_CUDA(cudaStreamSynchronize(cx.st[l]));
  return 0;
}


// ---- rsb_jet_render_static: the same renderer for callers that must not allocate ----
// Who needs it: a host that hands this library a fixed device-memory budget at start-up
// and forbids any later cudaMalloc -- e.g. a data-loader plugin living inside a training
// process whose framework owns (and has already reserved) the whole device through its own
// caching allocator, or a service whose memory is accounted per component.  For such a
// caller rsb_jet_render's `workspace_dev` is a problem when the image does not fit in
// shared memory (size >= ~236): it would have to carve the scratch out of somebody else's
// pool.  Here the raw-image slots live in the library image itself instead.
//
// The array below is reserved on EVERY GPU that loads librsb200.so, from load until the
// process exits, whether or not this entry point is ever called.
constexpr size_t kStaticScratchBytes = (size_t)32 << 20;  // 32 MiB
constexpr int kStaticMaxSize = 512;                        // largest image side served
static __device__ __align__(256) float g_jet_static_scratch[kStaticScratchBytes / sizeof(float)];

namespace {
std::mutex g_static_mu;
float* g_static_ptr[64] = {};  // per device: device address of g_jet_static_scratch
}  // namespace

int rsb_jet_render_static(const double* params, const int32_t* zoom_pad, int n, int slots, int size,
                          unsigned flags, void* out_img, int32_t* out_margin, double* out_peak,
                          void* stream) {
  if (int e = check_common("rsb_jet_render_static", n, slots, size)) return e;
  RSB_ARG(size <= kStaticMaxSize,
          "rsb_jet_render_static: size %d exceeds the built-in scratch limit of %d "
          "(use rsb_jet_render with a caller-provided workspace)", size, kStaticMaxSize);
  if (n == 0) return 0;
  RSB_ARG(params && zoom_pad && out_img && out_margin && out_peak, "rsb_jet_render_static: NULL buffer");
  RSB_ARG(((uintptr_t)params & 15) == 0, "rsb_jet_render_static: params must be 16-byte aligned");
  int dev = 0;
  RSB_CUDA(cudaGetDevice(&dev));
  RSB_ARG(dev >= 0 && dev < 64, "rsb_jet_render_static: device index %d out of range", dev);
  float* scratch = nullptr;
  {
    // One-time set-up per device, done by the first call on that device (which is why
    // that call must not happen while a stream is being captured).  In host code the
    // name g_jet_static_scratch is only a stand-in; the kernel needs the device address.
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
      float* ptr = nullptr;
      RSB_CUDA(cudaGetSymbolAddress((void**)&ptr, g_jet_static_scratch));
      RSB_ARG(ptr != nullptr, "rsb_jet_render_static: no device address for the scratch array");
      g_static_ptr[dev] = ptr;
    }
    scratch = g_static_ptr[dev];
  }
  {
    // sizes the on-chip renderer takes (up to 256 pixels) need no scratch at all: same kernel as
    // rsb_jet_render, nothing allocated, the built-in array untouched
    int cluster = 1;
    size_t smem = 0;
    if (v2_supported(size, slots, &cluster, &smem)) {
      Args a2;
      a2.params = params; a2.zoom_pad = zoom_pad; a2.n = n; a2.slots = slots; a2.size = size;
      a2.flags = flags; a2.out = out_img; a2.margin = out_margin; a2.peak = out_peak;
      a2.scratch = nullptr; a2.pitch = (size + 3) & ~3;
      return v2_launch(a2, cluster, smem, (cudaStream_t)stream);
    }
    if (bands_supported(size, slots, &smem))
      return bands_launch(params, zoom_pad, n, slots, size, flags, out_img, out_margin, out_peak, smem, stream);
  }
  Plan p;
  if (int e = make_plan(size, &p)) return e;
  int grid = n < p.grid_cap ? n : p.grid_cap;
  if (!p.smem_raw) {
    const size_t slot = (size_t)p.pitch * size * sizeof(float);
    const int fit = (int)(kStaticScratchBytes / slot);
    RSB_ARG(fit >= 1, "rsb_jet_render_static: one %dx%d image does not fit the %zu-byte scratch",
            size, size, kStaticScratchBytes);
    if (grid > fit) grid = fit;  // fewer resident CTAs, same results (one slot per CTA)
  }
  // Nothing has to start at zero: every raw pixel of a slot is written (phase 1) before
  // anything reads it, so whatever the previous call left behind is never observed.
  Args a;
  a.params = params; a.zoom_pad = zoom_pad; a.n = n; a.slots = slots; a.size = size;
  a.flags = flags; a.out = out_img; a.margin = out_margin; a.peak = out_peak;
  a.scratch = scratch; a.pitch = p.pitch;
  return launch(a, p, grid, (cudaStream_t)stream);
}

}  // extern "C"
