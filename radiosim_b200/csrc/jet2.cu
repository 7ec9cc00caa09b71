// On-chip jet renderer: the raw image never leaves shared memory.
//
// One image per thread-block CLUSTER.  kCluster == 1: the whole raw image sits in one
// CTA's shared memory (size <= ~224).  kCluster == 2: the image is split into two row
// bands, one per CTA of a 2-CTA cluster (size 256: 129 x 256 floats = 132 KB each);
// the composite zoom operator is centre preserving, so each CTA resamples its own band
// and the only data that crosses CTAs is ONE halo row, the margin and the arg-max,
// written straight into the peer's shared memory (DSMEM) before a cluster barrier.
//
// Per image (see jet.cu for the algorithm and reference citations):
//   setup   : 1-D bulk async copy of the parameter block (double buffered), warp 0
//             builds conic coefficients + bounding boxes.
//   phase 1 : warps pull 4-row x 128-column tiles from a shared counter; 16 pixels per
//             lane, ~6.5 issue slots per MUFU.EX2, components culled per tile.
//   zoom    : margin -> tap table -> H pass in place (column taps live in registers) ->
//             halo row exchange -> V pass (max) -> V pass (scale + single write).
#include <cooperative_groups.h>

#include "jet_common.cuh"

namespace cg = cooperative_groups;

namespace rsb {
namespace jet {
namespace v2 {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kTileRows = 4;

struct __align__(16) Fixed2 {
  double params[2][kMaxSlots * RSB_JET_PARAM_STRIDE];
  Coef coef[kMaxSlots];
  double pix64[16 * kMaxSlots];  // float64 slot terms of the <= 16 raw pixels under the peak
  uint64_t bar[2];
  unsigned long long xbest[2][2];  // [image parity][cluster rank] arg-max keys
  unsigned long long red_u[kWarps];
  double peak;
  int xmargin[2][2];             // [image parity][cluster rank] margins
  int red_i[kWarps];
  int nlive;
  int next_tile;
  int foot_lo, foot_hi;          // output coordinates with a non-zero tap
  float delta;
};

// one slot's float64 value at a raw pixel, operation order of gauss.py:77-88
__device__ inline double gauss64(const double* p, int r, int c) {
  const double amp = p[0];
  if (amp == 0.0) return 0.0;
  const double cr = cos(p[5]), sr = sin(p[5]);
  const double ra = p[2] * cr - p[1] * sr;
  const double rb = p[2] * sr + p[1] * cr;
  const double ua = r * cr - c * sr;
  const double ub = r * sr + c * cr;
  const double a = (ra - ua) / p[4];
  const double b = (rb - ub) / p[3];
  return amp * exp(-(a * a + b * b) / 2.0);
}

__device__ __forceinline__ unsigned long long best_key(float v, int idx) {
  return ((unsigned long long)__float_as_uint(fmaxf(v, 0.f)) << 32) | (unsigned)(~(unsigned)idx);
}

__device__ __forceinline__ void store_out4(void* out_row, bool f64, int c0, int N, bool vec,
                                           float4 v) {
  if (!f64) {
    float* row = reinterpret_cast<float*>(out_row);
    if (vec) {
      *reinterpret_cast<float4*>(row + c0) = v;
    } else {
      const float a[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (c0 + j < N) row[c0 + j] = a[j];
    }
  } else {
    double* row = reinterpret_cast<double*>(out_row);
    if (vec) {
      double2* p = reinterpret_cast<double2*>(row + c0);
      p[0] = make_double2((double)v.x, (double)v.y);
      p[1] = make_double2((double)v.z, (double)v.w);
    } else {
      const float a[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (c0 + j < N) row[c0 + j] = (double)a[j];
    }
  }
}

// local storage row of a V-pass tap; only zero-weight taps can fall outside [0, H]
__device__ __forceinline__ int lrow(int l, int H) { return min(max(l, 0), H); }

// column taps of 4 consecutive output columns, kept in registers during the H pass
struct ColTaps {
  float w[4][4];  // [tap][column]
  int ia[4];      // a | a1 << 16   (taps 0, 1)
  int ib[4];      // b | b1 << 16   (taps 2, 3)
};

template <int kCluster, int NS>
__global__ void __launch_bounds__(kThreads, 1) jet2_kernel(const Args a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Fixed2& sm = *reinterpret_cast<Fixed2*>(smem_raw);
  const int N = a.size, pitch = a.pitch, slots = a.slots;
  const int npad = (N + 3) & ~3;
  float* tab_w = reinterpret_cast<float*>(smem_raw + sizeof(Fixed2));  // [4][npad]
  int* tab_a = reinterpret_cast<int*>(tab_w + 4 * npad);               // [npad] taps 0,1
  int* tab_b = tab_a + npad;                                           // [npad] taps 2,3
  float* raw = reinterpret_cast<float*>(tab_b + npad);  // [H (+1 halo)][pitch]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int rank = (kCluster > 1) ? (int)(blockIdx.x % kCluster) : 0;
  const int cid = blockIdx.x / kCluster, nclusters = gridDim.x / kCluster;
  const int H = N / kCluster;             // rows of this CTA's band (N even when kCluster == 2)
  const int R0 = rank * H;                // first absolute row of the band
  const int base = (rank == 0) ? 0 : H - 1;  // absolute row stored at raw row 0 (rank 1: halo)
  Fixed2* peer = nullptr;
  float* peer_raw = nullptr;
  if constexpr (kCluster > 1) {
    cg::cluster_group cluster = cg::this_cluster();
    peer = cluster.map_shared_rank(&sm, rank ^ 1);
    peer_raw = cluster.map_shared_rank(raw, rank ^ 1);
  }

  const uint32_t pbytes = (uint32_t)slots * RSB_JET_PARAM_STRIDE * 8u;
  const int strips = (N + kStrip - 1) / kStrip;
  const bool vec = (N & 3) == 0;
  const bool f64 = (a.flags & RSB_JET_OUT_F64) != 0;
  const bool normalize = (a.flags & RSB_JET_NORMALIZE) != 0;
  const size_t esz = f64 ? 8 : 4;

  if (tid == 0) {
    mbar_init(&sm.bar[0], 1);
    mbar_init(&sm.bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();
  if (tid == 0 && cid < a.n) {
    mbar_expect_tx(&sm.bar[0], pbytes);
    bulk_g2s(sm.params[0], a.params + (size_t)cid * slots * 6, pbytes, &sm.bar[0]);
  }
  if constexpr (kCluster > 1) cg::this_cluster().sync();  // peer CTA is resident

  int it = 0;
  for (int img = cid; img < a.n; img += nclusters, ++it) {
    const int buf = it & 1;
    if (tid == 0) {
      sm.next_tile = 0;
      if (img + nclusters < a.n) {
        fence_proxy_async();
        mbar_expect_tx(&sm.bar[buf ^ 1], pbytes);
        bulk_g2s(sm.params[buf ^ 1], a.params + (size_t)(img + nclusters) * slots * 6, pbytes,
                 &sm.bar[buf ^ 1]);
      }
    }
    mbar_wait(&sm.bar[buf], (it >> 1) & 1);
    const double* P = sm.params[buf];
    const int zpad = a.zoom_pad[img];
    const bool zoom = zpad >= 0;
    if (warp == 0) build_coefs(sm, P, slots, N, lane);
    __syncthreads();
    const int nlive = sm.nlive;
    const float thr = (float)kEdgeThr, delta = sm.delta;

    // ---------------- phase 1: splat this CTA's band, 4 x 128 tiles ------------------
    int my_margin = N;
    unsigned long long my_best = 0ull;
    float best_v = -1.f;
    const int ntiles = ((H + kTileRows - 1) / kTileRows) * strips;
    for (;;) {
      int t = 0;
      if (lane == 0) t = atomicAdd(&sm.next_tile, 1);
      t = __shfl_sync(0xffffffffu, t, 0);
      if (t >= ntiles) break;
      const int tr = t / strips, s = t - tr * strips;
      const int r0 = R0 + tr * kTileRows, sc0 = s * kStrip, c0 = sc0 + 4 * lane;
      const float r0f = (float)r0, c0f = (float)c0;
      float acc[kTileRows][4];
#pragma unroll
      for (int i = 0; i < kTileRows; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
      for (int k = 0; k < nlive; ++k) {
        const int4 bb = *reinterpret_cast<const int4*>(&sm.coef[k].r0);
        if (r0 > bb.y || r0 + kTileRows - 1 < bb.x || sc0 > bb.w || sc0 + kStrip - 1 < bb.z)
          continue;  // warp-uniform cull
        const float4 q = *reinterpret_cast<const float4*>(&sm.coef[k].A);    // A B C amp
        const float4 ct = *reinterpret_cast<const float4*>(&sm.coef[k].irf);  // irf fr icf fc
        const float dc = (c0f - ct.z) - ct.w;
        float Pj[4], Qj[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float d = dc + (float)j;
          Qj[j] = q.y * d;
          Pj[j] = q.x * d * d;
        }
#pragma unroll
        for (int i = 0; i < kTileRows; ++i) {
          const float dr = ((r0f + (float)i) - ct.x) - ct.y;
          const float Cq = q.z * dr * dr;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            acc[i][j] = fmaf(q.w, ex2(fmaf(dr, Qj[j], Pj[j] + Cq)), acc[i][j]);
        }
      }
      if (c0 < N) {
#pragma unroll
        for (int i = 0; i < kTileRows; ++i) {
          const int r = r0 + i;
          if (r >= R0 + H) break;
          *reinterpret_cast<float4*>(raw + (size_t)(r - base) * pitch + c0) =
              make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
          const float top = fmaxf(fmaxf(acc[i][0], acc[i][1]), fmaxf(acc[i][2], acc[i][3]));
          if (zoom) {
            if (top > thr - delta) {
              const int db = min(r, N - 1 - r);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int c = c0 + j;
                const int d = min(db, min(c, N - 1 - c));
                if (c < N && d < my_margin && acc[i][j] > thr - delta) {
                  bool above = acc[i][j] > thr;
                  if (acc[i][j] < thr + delta) above = eval64(P, slots, r, c) > kEdgeThr;
                  if (above) my_margin = d;
                }
              }
            }
          } else if (normalize && top >= best_v) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (c0 + j < N) {
                const unsigned long long key = best_key(acc[i][j], r * N + c0 + j);
                my_best = key > my_best ? key : my_best;
              }
            }
            best_v = __uint_as_float((unsigned)(my_best >> 32));
          }
        }
      }
    }

    double peak = 1.0;
    int margin = 0;
    char* out = reinterpret_cast<char*>(a.out) + (size_t)img * N * N * esz;

    if (!zoom) {
      // ------------- plain branch ---------------------------------------------------
      if (normalize) {
        my_best = warp_max_u64(my_best);
        if (lane == 0) sm.red_u[warp] = my_best;
        __syncthreads();
        if (warp == 0) {
          unsigned long long b = lane < kWarps ? sm.red_u[lane] : 0ull;
          b = warp_max_u64(b);
          if (lane == 0) {
            sm.xbest[buf][rank] = b;
            if constexpr (kCluster > 1) peer->xbest[buf][rank] = b;
          }
        }
        if constexpr (kCluster > 1) cg::this_cluster().sync(); else __syncthreads();
        unsigned long long b = sm.xbest[buf][0];
        if constexpr (kCluster > 1) b = sm.xbest[buf][1] > b ? sm.xbest[buf][1] : b;
        const unsigned idx = ~(unsigned)(b & 0xffffffffull);
        const int pr = (int)(idx / (unsigned)N), pc = (int)(idx % (unsigned)N);
        for (int k = tid; k < slots; k += kThreads) sm.pix64[k] = gauss64(P + 6 * k, pr, pc);
        __syncthreads();
        if (tid == 0) {
          double s = 0.0;
          for (int k = 0; k < slots; ++k) s += sm.pix64[k];
          sm.peak = s;
        }
        __syncthreads();
        peak = sm.peak;
      } else {
        __syncthreads();
      }
      const float inv = normalize ? (float)(1.0 / peak) : 1.f;
      const int q4 = npad >> 2;
      for (int e = tid; e < H * q4; e += kThreads) {
        const int rl = e / q4, c0 = (e - rl * q4) << 2;
        const int r = R0 + rl;
        float4 v = *reinterpret_cast<const float4*>(raw + (size_t)(r - base) * pitch + c0);
        v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
        store_out4(out + (size_t)r * N * esz, f64, c0, N, vec, v);
      }
    } else {
      // ------------- zoom branch ------------------------------------------------------
      my_margin = warp_min(my_margin);
      if (lane == 0) sm.red_i[warp] = my_margin;
      __syncthreads();
      if (warp == 0) {
        int m = lane < kWarps ? sm.red_i[lane] : N;
        m = warp_min(m);
        if (lane == 0) {
          sm.xmargin[buf][rank] = m;
          if constexpr (kCluster > 1) peer->xmargin[buf][rank] = m;
          sm.foot_lo = N;
          sm.foot_hi = -1;
        }
      }
      if constexpr (kCluster > 1) cg::this_cluster().sync(); else __syncthreads();
      margin = sm.xmargin[buf][0];
      if constexpr (kCluster > 1) margin = min(margin, sm.xmargin[buf][1]);
      if (margin >= N) margin = 0;  // empty mask -> idx 0 (argmax of an all-false array)

      // composite tap table (float64 math, float32 weights) + output footprint
      {
        int lo = N, hi = -1;
        for (int d = tid; d < npad; d += kThreads) {
          float w[4] = {0.f, 0.f, 0.f, 0.f};
          int ia = margin | (margin << 16), ib = ia;
          if (d < N) {
            const Taps64 t = composite_taps(d, N, margin, zpad);
#pragma unroll
            for (int k = 0; k < 4; ++k) w[k] = (float)t.w[k];
            ia = t.i[0] | (t.i[1] << 16);
            ib = t.i[2] | (t.i[3] << 16);
            if (w[0] != 0.f || w[1] != 0.f || w[2] != 0.f || w[3] != 0.f) {
              lo = min(lo, d);
              hi = max(hi, d);
            }
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) tab_w[k * npad + d] = w[k];
          tab_a[d] = ia;
          tab_b[d] = ib;
        }
        lo = warp_min(lo);
        hi = -warp_min(-hi);
        if (lane == 0) {
          if (lo < N) atomicMin(&sm.foot_lo, lo);
          if (hi >= 0) atomicMax(&sm.foot_hi, hi);
        }
      }
      __syncthreads();
      const int flo = sm.foot_lo, fhi = sm.foot_hi;  // empty footprint: flo = N, fhi = -1

      // H pass, in place: row r of raw becomes row r of T.  Column taps stay in registers.
      {
        ColTaps ct[NS];
        bool act[NS];
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          const int sc0 = s * kStrip, c0 = sc0 + 4 * lane;
          act[s] = (s < strips) && sc0 <= fhi && sc0 + kStrip - 1 >= flo;
          if (act[s] && c0 < N) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float4 w = *reinterpret_cast<const float4*>(tab_w + k * npad + c0);
              ct[s].w[k][0] = w.x; ct[s].w[k][1] = w.y; ct[s].w[k][2] = w.z; ct[s].w[k][3] = w.w;
            }
            const int4 ia = *reinterpret_cast<const int4*>(tab_a + c0);
            const int4 ib = *reinterpret_cast<const int4*>(tab_b + c0);
            ct[s].ia[0] = ia.x; ct[s].ia[1] = ia.y; ct[s].ia[2] = ia.z; ct[s].ia[3] = ia.w;
            ct[s].ib[0] = ib.x; ct[s].ib[1] = ib.y; ct[s].ib[2] = ib.z; ct[s].ib[3] = ib.w;
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
#pragma unroll
              for (int k = 0; k < 4; ++k) ct[s].w[k][j] = 0.f;
              ct[s].ia[j] = ct[s].ib[j] = margin | (margin << 16);
            }
          }
        }
        const int rlo = max(R0, margin), rhi = min(R0 + H, N - margin);
        for (int r = rlo + warp; r < rhi; r += kWarps) {
          float* row = raw + (size_t)(r - base) * pitch;
          float4 res[NS];
#pragma unroll
          for (int s = 0; s < NS; ++s) {
            float o[4] = {0.f, 0.f, 0.f, 0.f};
            if (act[s]) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int ia = ct[s].ia[j], ib = ct[s].ib[j];
                o[j] = ct[s].w[0][j] * row[ia & 0xffff];
                o[j] = fmaf(ct[s].w[1][j], row[ia >> 16], o[j]);
                o[j] = fmaf(ct[s].w[2][j], row[ib & 0xffff], o[j]);
                o[j] = fmaf(ct[s].w[3][j], row[ib >> 16], o[j]);
              }
            }
            res[s] = make_float4(o[0], o[1], o[2], o[3]);
          }
          __syncwarp();
#pragma unroll
          for (int s = 0; s < NS; ++s) {
            const int c0 = s * kStrip + 4 * lane;
            if (act[s] && c0 < N) *reinterpret_cast<float4*>(row + c0) = res[s];
          }
        }
      }
      __syncthreads();
      if constexpr (kCluster > 1) {
        // halo: my boundary row of T goes into the peer's halo slot
        const int src = (rank == 0) ? (H - 1) : 1;   // local row index of abs row H-1 / H
        const int dst = (rank == 0) ? 0 : H;         // peer's halo slot
        for (int e = tid; e < (pitch >> 2); e += kThreads)
          reinterpret_cast<float4*>(peer_raw + (size_t)dst * pitch)[e] =
              reinterpret_cast<const float4*>(raw + (size_t)src * pitch)[e];
        cg::this_cluster().sync();
      }

      // V pass over the footprint rows of this band
      const int vr0 = max(R0, flo), vr1 = min(R0 + H - 1, fhi);   // inclusive
      const int vs0 = flo < N ? flo / kStrip : 0, vs1 = fhi >= 0 ? fhi / kStrip : -1;
      const int vrows = max(vr1 - vr0 + 1, 0), vstr = max(vs1 - vs0 + 1, 0);
      float inv = 1.f;
      if (normalize) {
        my_best = 0ull;
        best_v = -1.f;
        for (int item = warp; item < vrows * vstr; item += kWarps) {
          const int r = vr0 + item / vstr, s = vs0 + item % vstr;
          const int c0 = s * kStrip + 4 * lane;
          if (c0 >= N) continue;
          const float w0 = tab_w[r], w1 = tab_w[npad + r], w2 = tab_w[2 * npad + r],
                      w3 = tab_w[3 * npad + r];
          const int ia = tab_a[r], ib = tab_b[r];
          const float4 t0 = *reinterpret_cast<const float4*>(raw + (size_t)lrow((ia & 0xffff) - base, H) * pitch + c0);
          const float4 t1 = *reinterpret_cast<const float4*>(raw + (size_t)lrow((ia >> 16) - base, H) * pitch + c0);
          const float4 t2 = *reinterpret_cast<const float4*>(raw + (size_t)lrow((ib & 0xffff) - base, H) * pitch + c0);
          const float4 t3 = *reinterpret_cast<const float4*>(raw + (size_t)lrow((ib >> 16) - base, H) * pitch + c0);
          const float v[4] = {fmaf(w3, t3.x, fmaf(w2, t2.x, fmaf(w1, t1.x, w0 * t0.x))),
                              fmaf(w3, t3.y, fmaf(w2, t2.y, fmaf(w1, t1.y, w0 * t0.y))),
                              fmaf(w3, t3.z, fmaf(w2, t2.z, fmaf(w1, t1.z, w0 * t0.z))),
                              fmaf(w3, t3.w, fmaf(w2, t2.w, fmaf(w1, t1.w, w0 * t0.w)))};
          const float top = fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3]));
          if (top >= best_v) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (c0 + j < N) {
                const unsigned long long key = best_key(v[j], r * N + c0 + j);
                my_best = key > my_best ? key : my_best;
              }
            }
            best_v = __uint_as_float((unsigned)(my_best >> 32));
          }
        }
        my_best = warp_max_u64(my_best);
        if (lane == 0) sm.red_u[warp] = my_best;
        __syncthreads();
        if (warp == 0) {
          unsigned long long b = lane < kWarps ? sm.red_u[lane] : 0ull;
          b = warp_max_u64(b);
          if (lane == 0) {
            sm.xbest[buf][rank] = b;
            if constexpr (kCluster > 1) peer->xbest[buf][rank] = b;
          }
        }
        if constexpr (kCluster > 1) cg::this_cluster().sync(); else __syncthreads();
        unsigned long long b = sm.xbest[buf][0];
        if constexpr (kCluster > 1) b = sm.xbest[buf][1] > b ? sm.xbest[buf][1] : b;
        // float64 value of the arg-max pixel: <= 4x4 raw pixels x slots, float64 weights
        const unsigned idx = ~(unsigned)(b & 0xffffffffull);
        const int pr = (int)(idx / (unsigned)N), pc = (int)(idx % (unsigned)N);
        const Taps64 tr = composite_taps(pr, N, margin, zpad);
        const Taps64 tc = composite_taps(pc, N, margin, zpad);
        for (int e = tid; e < 16 * slots; e += kThreads) {
          const int px = e / slots, k = e - px * slots;
          const int i = px >> 2, j = px & 3;
          sm.pix64[e] = (tr.w[i] * tc.w[j] != 0.0) ? gauss64(P + 6 * k, tr.i[i], tc.i[j]) : 0.0;
        }
        __syncthreads();
        if (warp == 0) {
          double s = 0.0;
          if (lane < 16)
            for (int k = 0; k < slots; ++k) s += sm.pix64[lane * slots + k];
          double accd = 0.0;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            double h = 0.0;
#pragma unroll
            for (int j = 0; j < 4; ++j) h += tc.w[j] * __shfl_sync(0xffffffffu, s, i * 4 + j);
            accd += tr.w[i] * h;
          }
          if (lane == 0) sm.peak = accd;
        }
        __syncthreads();
        peak = sm.peak;
        inv = (float)(1.0 / peak);
      }

      // V pass B: every output pixel of the band written once, scaled
      for (int item = warp; item < H * strips; item += kWarps) {
        const int r = R0 + item / strips, s = item % strips;
        const int c0 = s * kStrip + 4 * lane;
        if (c0 >= N) continue;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r >= vr0 && r <= vr1 && s >= vs0 && s <= vs1) {
          const float w0 = tab_w[r] * inv, w1 = tab_w[npad + r] * inv, w2 = tab_w[2 * npad + r] * inv,
                      w3 = tab_w[3 * npad + r] * inv;
          const int ia = tab_a[r], ib = tab_b[r];
          const float4 t0 = *reinterpret_cast<const float4*>(raw + (size_t)lrow((ia & 0xffff) - base, H) * pitch + c0);
          const float4 t1 = *reinterpret_cast<const float4*>(raw + (size_t)lrow((ia >> 16) - base, H) * pitch + c0);
          const float4 t2 = *reinterpret_cast<const float4*>(raw + (size_t)lrow((ib & 0xffff) - base, H) * pitch + c0);
          const float4 t3 = *reinterpret_cast<const float4*>(raw + (size_t)lrow((ib >> 16) - base, H) * pitch + c0);
          v.x = fmaf(w3, t3.x, fmaf(w2, t2.x, fmaf(w1, t1.x, w0 * t0.x)));
          v.y = fmaf(w3, t3.y, fmaf(w2, t2.y, fmaf(w1, t1.y, w0 * t0.y)));
          v.z = fmaf(w3, t3.z, fmaf(w2, t2.z, fmaf(w1, t1.z, w0 * t0.z)));
          v.w = fmaf(w3, t3.w, fmaf(w2, t2.w, fmaf(w1, t1.w, w0 * t0.w)));
        }
        store_out4(out + (size_t)r * N * esz, f64, c0, N, vec, v);
      }
    }
    if (tid == 0 && rank == 0) {
      a.margin[img] = margin;
      a.peak[img] = peak;
    }
    __syncthreads();  // raw / tab / coef / params[buf] are reused by the next image
  }
  if constexpr (kCluster > 1) cg::this_cluster().sync();  // no CTA exits while its smem is in use
}


// ------------------------------------------------------------------ host side
template <int kCluster, int NS>
static int launch_v2(const Args& a, size_t smem_bytes, cudaStream_t st) {
  auto k = jet2_kernel<kCluster, NS>;
  RSB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  const int sms = sm_count();
  int nclusters = sms / kCluster;
  if (nclusters > a.n) nclusters = a.n;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(nclusters * kCluster));
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  RSB_CUDA(cudaLaunchKernelEx(&cfg, k, a));
  RSB_LAUNCHED();
  return 0;
}

}  // namespace v2

static size_t v2_bytes(int size, int cluster) {
  const int npad = (size + 3) & ~3;
  const int rows = size / cluster + (cluster > 1 ? 1 : 0);
  return sizeof(v2::Fixed2) + (size_t)npad * 24 + (size_t)rows * npad * 4;
}

bool v2_supported(int size, int* cluster, size_t* smem_bytes) {
  int dev = 0, max_optin = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess)
    return false;
  if ((size + kStrip - 1) / kStrip > 3) return false;
  if (v2_bytes(size, 1) <= (size_t)max_optin) {
    *cluster = 1;
    *smem_bytes = v2_bytes(size, 1);
    return true;
  }
  if ((size & 1) == 0 && v2_bytes(size, 2) <= (size_t)max_optin) {
    *cluster = 2;
    *smem_bytes = v2_bytes(size, 2);
    return true;
  }
  return false;
}

int v2_launch(const Args& a, int cluster, size_t smem_bytes, cudaStream_t st) {
  const int strips = (a.size + kStrip - 1) / kStrip;
  if (cluster == 1) {
    switch (strips) {
      case 1: return v2::launch_v2<1, 1>(a, smem_bytes, st);
      case 2: return v2::launch_v2<1, 2>(a, smem_bytes, st);
      default: return v2::launch_v2<1, 3>(a, smem_bytes, st);
    }
  }
  switch (strips) {
    case 1: return v2::launch_v2<2, 1>(a, smem_bytes, st);
    case 2: return v2::launch_v2<2, 2>(a, smem_bytes, st);
    default: return v2::launch_v2<2, 3>(a, smem_bytes, st);
  }
}

}  // namespace jet
}  // namespace rsb
