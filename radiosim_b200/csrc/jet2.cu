// On-chip jet renderer: the raw image never leaves shared memory.
//
// One image per thread-block CLUSTER.  kCluster == 1: the whole raw image sits in one
// CTA's shared memory (size <= 128).  kCluster == 2: the image is split into two row
// bands, one per CTA of a 2-CTA cluster (size 256: 129 x 256 floats = 132 KB each);
// the composite zoom operator is centre preserving, so each CTA resamples its own band
// and the only data that crosses CTAs is ONE halo row, the margin and the arg-max,
// written straight into the peer's shared memory (DSMEM) before a cluster barrier.
//
// Per image (see jet.cu for the algorithm and the reference citations):
//   setup   : 1-D bulk async copy of the parameter block (double buffered).  Warp 0
//             builds, one image AHEAD and off the critical path, the float32 conic
//             coefficients and, per component and 8-row group, the exact column
//             interval of its > tau footprint (the culling table).
//   phase 1 : warps pull 8-row x 64-column tiles from a shared counter; 16 pixels per
//             lane, ~6.4 issue slots per MUFU.EX2; a component is skipped when its
//             column interval for the tile's row group misses the tile.
//   zoom    : margin (DSMEM atomicMin) -> tap table -> H pass in place, rows pulled
//             dynamically, column taps in registers -> halo row exchange -> ONE V pass
//             whose results stay in registers -> arg-max (DSMEM atomicMax) -> scaled
//             128-bit stores.  The float64 peak for the labels is evaluated by one
//             warp while the others already render the next image.
#include <cooperative_groups.h>

#include "jet_common.cuh"

namespace cg = cooperative_groups;

namespace rsb {
namespace jet {
namespace v2 {

#ifdef RSB_TIMELINE
__device__ long long* g_timeline = nullptr;  // [image][rank][16] clock64 stamps (debug builds)
#define RSB_STAMP_T(k, t) do { if (g_timeline && tid == (t)) g_timeline[((size_t)img * 2 + rank) * 16 + (k)] = clock64(); } while (0)
#define RSB_STAMP(k) do { if (g_timeline && tid == 64) g_timeline[((size_t)img * 2 + rank) * 16 + (k)] = clock64(); } while (0)
#else
#define RSB_STAMP(k) do {} while (0)
#define RSB_STAMP_T(k, t) do {} while (0)
#endif

constexpr int kWorkers = 512;             // threads that render
constexpr int kWarps = kWorkers / 32;
constexpr int kThreads = kWorkers;
constexpr int kTileRows = 8;
constexpr int kTileCols = 64;   // 2 columns per lane
constexpr int kMaxBand = 128;   // rows per CTA band (V-pass results live in registers)
constexpr int kRowsPerWarp = kMaxBand / kWarps;
constexpr int kMaxGroups = kMaxBand / kTileRows;
constexpr int kMaxCand = 64;
constexpr int kMaxTiles = kMaxGroups * 4;  // 64-column tiles of a band (size <= 256)

struct __align__(16) Fixed2 {
  double params[3][kMaxSlots * RSB_JET_PARAM_STRIDE];  // triple buffered (bulk async copies)
  Coef coef[2][kMaxSlots];                             // double buffered, built one image ahead
  int span[2][kMaxSlots][kMaxGroups];  // lo | hi << 16 column interval per (live slot, row group)
  double pix64[20 * kMaxSlots];     // peak64 scratch: 16 pixel x slot terms + 4 per-slot rotations
  uint64_t bar[3];
  unsigned long long xval[3][2];    // peer's margin / (unused) / arg-max key, [kind][use parity]
  uint64_t xbar[3];                 // mbarriers the peer's st.async completes: margin, halo, best
  unsigned long long lbest;         // this CTA's running arg-max key
  int lmargin;                      // this CTA's running margin
  int nlive[2];
  int next_tile;
  int next_row;
  int foot_lo, foot_hi;             // output coordinates with a non-zero tap
  float delta[2];
  int dlo[2];                       // no pixel closer than this to the border exceeds 0.01
  struct LabelReq { unsigned long long best; int img, pb, margin, zpad; } label[2];  // [image parity]
  float geo[kMaxSlots * 8];         // build_coefs2 scratch (SlotGeo per live slot)
  int ncand;                        // borderline pixels awaiting a float64 verdict
  int cand[kMaxCand];               // r << 16 | c
  unsigned char tile_order[kMaxTiles];  // tiles sorted by distance from the image centre
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

__device__ __forceinline__ void store_out4(char* out_row, bool f64, int c0, int N, bool vec,
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

// ---- one warp: parameter block -> compacted coefficients + culling spans + dlo ---------
// Runs while the other warps render the previous image, so it is written for latency:
// float32 throughout (only the centre split needs the float64 parameters) and every lane
// busy.  Ellipse of the > tau footprint: z^T M z <= T^2 with z = (dr, dc).  For a row range
// the covered columns are [min cl, max cr], cl/cr = centre line -/+ half width; cr is
// concave in dr, so its maximum over the range is the ellipse's rightmost point if that
// row lies in the range, else the larger end value (mirror image for cl).
struct SlotGeo {  // per live slot, shared between the lanes through shared memory
  float Mrr, Mcc, Mrc, T2, ec, x, y;
  int degenerate;
};

__device__ inline void build_coefs2(Fixed2& sm, int cb, const double* P, int slots, int N, int R0,
                                    int H, int lane, bool zoom) {
  SlotGeo* geo = reinterpret_cast<SlotGeo*>(sm.geo);
  float amax = 0.f;
  for (int k = lane; k < slots; k += 32) amax = fmaxf(amax, fabsf((float)P[k * 6]));
  amax = warp_max(amax);
  const float tau = 2e-8f * amax;  // <= 19 x 2e-8 of the peak is dropped in total
  const int ngroups = (H + kTileRows - 1) / kTileRows;
  int base = 0;
  // frame bound (dlo) inputs of this lane's slots: k = lane, lane + 32
  float ka[2] = {0.f, 0.f}, kx[2] = {0.f, 0.f}, ky[2] = {0.f, 0.f}, cx[2] = {0.f, 0.f}, cy[2] = {0.f, 0.f};
  bool certified = false, all_nonneg = true;
  for (int k0 = 0, h = 0; k0 < slots; k0 += 32, ++h) {
    const int k = k0 + lane;
    bool live = false;
    Coef c;
    SlotGeo g;
    g.degenerate = 0;
    if (k < slots) {
      const double* p = P + k * 6;
      const double xd = p[1], yd = p[2];
      const float amp = (float)p[0], sx = (float)p[3], sy = (float)p[4];
      if (p[0] != 0.0 && fabsf(amp) > tau) {
        float sr, cr;
        sincosf((float)p[5], &sr, &cr);
        const float iu = 1.f / (sy * sy), iv = 1.f / (sx * sx);
        g.Mrr = cr * cr * iu + sr * sr * iv;
        g.Mcc = sr * sr * iu + cr * cr * iv;
        g.Mrc = sr * cr * (iv - iu);
        const float hl = 0.72134752f;  // log2(e) / 2
        c.A = -hl * g.Mcc;
        c.C = -hl * g.Mrr;
        c.B = -2.f * hl * g.Mrc;
        c.amp = amp;
        const double yr = rint(yd), xr = rint(xd);
        c.irf = (float)yr;
        c.fr = (float)(yd - yr);
        c.icf = (float)xr;
        c.fc = (float)(xd - xr);
        g.x = (float)xd;
        g.y = (float)yd;
        g.T2 = 2.f * logf(fabsf(amp) / tau);
        const float det = g.Mrr * g.Mcc - g.Mrc * g.Mrc;
        const float er = sqrtf(g.T2 * g.Mcc / det);
        g.ec = sqrtf(g.T2 * g.Mrr / det);
        g.degenerate = !(isfinite(er) && isfinite(g.ec));
        const float lo_r = fmaxf(floorf(g.y - er) - 1.f, 0.f), hi_r = fminf(ceilf(g.y + er) + 1.f, N - 1.f);
        const float lo_c = fmaxf(floorf(g.x - g.ec) - 1.f, 0.f), hi_c = fminf(ceilf(g.x + g.ec) + 1.f, N - 1.f);
        if (g.degenerate) {
          live = true;
          c.r0 = 0; c.r1 = N - 1; c.c0 = 0; c.c1 = N - 1;
        } else if (lo_r <= hi_r && lo_c <= hi_c) {
          live = true;
          c.r0 = (int)lo_r; c.r1 = (int)hi_r; c.c0 = (int)lo_c; c.c1 = (int)hi_c;
        }
        if (h < 2) {
          ka[h] = fabsf(amp) * 1.0001f;
          kx[h] = (det / g.Mrr) * 0.9999f * hl;
          ky[h] = (det / g.Mcc) * 0.9999f * hl;
          cx[h] = g.x;
          cy[h] = g.y;
          // the pixel nearest to the centre is at most half a pixel away in each coordinate:
          // this slot alone lifts it above the threshold -> the > 0.01 mask is not empty
          const float qmax = 0.25f * (g.Mrr + g.Mcc + 2.f * fabsf(g.Mrc));
          if (amp * __expf(-0.5f * qmax) > 0.0101f && g.x >= 0.f && g.x <= N - 1.f && g.y >= 0.f &&
              g.y <= N - 1.f)
            certified = true;
        }
      }
      if (p[0] < 0.0) all_nonneg = false;
      if (p[0] != 0.0 && h >= 2) all_nonneg = false;  // > 64 slots: no frame bound
    }
    const unsigned mask = __ballot_sync(0xffffffffu, live);
    if (live) {
      const int pos = base + __popc(mask & ((1u << lane) - 1u));
      sm.coef[cb][pos] = c;
      geo[pos] = g;
    }
    base += __popc(mask);
  }
  __syncwarp();
  // culling spans: (slot, row group) pairs dealt over the lanes
  for (int e = lane; e < base * ngroups; e += 32) {
    const int pos = e / ngroups, gi = e - pos * ngroups;
    const SlotGeo g = geo[pos];
    const Coef& c = sm.coef[cb][pos];
    const int ga = max(R0 + gi * kTileRows, c.r0);
    const int gb = min(min(R0 + gi * kTileRows + kTileRows - 1, R0 + H - 1), c.r1);
    int packed = 0x7fff;  // lo = 32767 > hi = 0: empty
    if (ga <= gb) {
      int lo_i = c.c0, hi_i = c.c1;
      if (!g.degenerate) {
        const float det = g.Mrr * g.Mcc - g.Mrc * g.Mrc;
        const float dr_right = -g.Mrc * g.ec / g.Mrr;  // row offset of the rightmost point
        const float inv_cc = 1.f / g.Mcc, slope = -g.Mrc * inv_cc;
        const float da = (float)ga - g.y, db = (float)gb - g.y;
        const float hwa = sqrtf(fmaxf(g.T2 * g.Mcc - det * da * da, 0.f)) * inv_cc;
        const float hwb = sqrtf(fmaxf(g.T2 * g.Mcc - det * db * db, 0.f)) * inv_cc;
        const float cla = slope * da, clb = slope * db;
        const float hi = (da <= dr_right && dr_right <= db) ? g.ec : fmaxf(cla + hwa, clb + hwb);
        const float lo = (da <= -dr_right && -dr_right <= db) ? -g.ec : fminf(cla - hwa, clb - hwb);
        lo_i = max((int)floorf(g.x + lo) - 1, c.c0);
        hi_i = min((int)ceilf(g.x + hi) + 1, c.c1);
      }
      if (lo_i <= hi_i) packed = lo_i | (hi_i << 16);
    }
    sm.span[cb][pos][gi] = packed;
  }
  // Zoom images only use the crop [margin, N-margin)^2 of the raw image, and margin >= dlo,
  // where dlo is the smallest border distance d at which an UPPER BOUND of the image over
  // the frame {border distance <= d} exceeds the 0.01 threshold.  Bound: sum over slots of
  // the slot's maximum over the frame = amp if its centre lies in the frame, else
  // amp * exp(-q/2) with q the smallest Mahalanobis distance to the four frame lines.
  int dlo = 0;
  if (zoom) {
    certified = __any_sync(0xffffffffu, certified) && __all_sync(0xffffffffu, all_nonneg);
    auto frame_bound = [&](int d) -> float {
      const float lo = (float)d, hi = (float)(N - 1 - d);
      float u = 0.f;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        if (ka[h] == 0.f) continue;
        float v = ka[h];
        if (!(cx[h] <= lo || cx[h] >= hi || cy[h] <= lo || cy[h] >= hi)) {
          const float dc0 = lo - cx[h], dc1 = hi - cx[h], dr0 = lo - cy[h], dr1 = hi - cy[h];
          const float q = fminf(fminf(dc0 * dc0, dc1 * dc1) * kx[h], fminf(dr0 * dr0, dr1 * dr1) * ky[h]);
          v *= ex2(-q);
        }
        u += v;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) u += __shfl_xor_sync(0xffffffffu, u, o);
      return u;
    };
    const float limit = 0.00999f;
    // without the certificate an empty mask (margin := 0, whole image used) cannot be ruled out
    if (certified && frame_bound(0) <= limit) {
      int a_ = 0, b_ = (N - 1) / 2;
      if (frame_bound(b_) <= limit) {
        a_ = -1;  // cannot happen with a certificate; stay safe
      } else {
        while (b_ - a_ > 1) {
          const int mid = (a_ + b_) >> 1;
          if (frame_bound(mid) <= limit) a_ = mid; else b_ = mid;
        }
      }
      dlo = a_ + 1;  // frames 0..a_ hold nothing above the threshold
    }
  }
  if (lane == 0) {
    sm.nlive[cb] = base;
    sm.delta[cb] = 2e-6f * fmaxf(1.f, amax);
    sm.dlo[cb] = dlo;
  }
}

// ---- float64 peak (jet_max) for the labels, one warp -------------------------------
// The float32 image fixes WHERE the maximum is; its float64 value is recomputed from the
// parameters: <= 4x4 raw pixels x slots (operation order of gauss.py:77-88), combined with
// float64 tap weights.  scratch: >= 16*slots + 4*slots doubles.
__device__ inline double peak64(double* scratch, const double* P, int slots, int N, bool zoom,
                                int zpad, int margin, unsigned long long image_best, int lane) {
  const unsigned idx = ~(unsigned)(image_best & 0xffffffffull);
  const int pr = (int)(idx / (unsigned)N), pc = (int)(idx % (unsigned)N);
  double* rot = scratch + 16 * slots;  // per slot: cr, sr, ra, rb
  for (int k = lane; k < slots; k += 32) {
    const double* p = P + 6 * k;
    const double cr = cos(p[5]), sr = sin(p[5]);
    rot[4 * k] = cr;
    rot[4 * k + 1] = sr;
    rot[4 * k + 2] = p[2] * cr - p[1] * sr;
    rot[4 * k + 3] = p[2] * sr + p[1] * cr;
  }
  __syncwarp();
  Taps64 tr, tc;
  int npx = 1;
  if (zoom) {
    tr = composite_taps(pr, N, margin, zpad);
    tc = composite_taps(pc, N, margin, zpad);
    npx = 16;
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      tr.i[k] = pr; tc.i[k] = pc;
      tr.w[k] = tc.w[k] = (k == 0) ? 1.0 : 0.0;
    }
  }
  for (int e = lane; e < npx * slots; e += 32) {
    const int px = e / slots, k = e - px * slots;
    const int i = px >> 2, j = px & 3;
    const double* p = P + 6 * k;
    double v = 0.0;
    if (p[0] != 0.0 && tr.w[i] * tc.w[j] != 0.0) {
      const int r = tr.i[i], c = tc.i[j];
      const double cr = rot[4 * k], sr = rot[4 * k + 1];
      const double ua = r * cr - c * sr, ub = r * sr + c * cr;
      const double a = (rot[4 * k + 2] - ua) / p[4], b = (rot[4 * k + 3] - ub) / p[3];
      v = p[0] * exp(-(a * a + b * b) / 2.0);
    }
    scratch[e] = v;
  }
  __syncwarp();
  double s = 0.0;
  if (lane < npx)
    for (int k = 0; k < slots; ++k) s += scratch[lane * slots + k];  // slot order
  double acc = 0.0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    double h = 0.0;
#pragma unroll
    for (int j = 0; j < 4; ++j) h += tc.w[j] * __shfl_sync(0xffffffffu, s, (i * 4 + j) & (npx - 1));
    acc += tr.w[i] * h;
  }
  __syncwarp();
  return acc;
}

// The composite operator has at most THREE consecutive taps: the two zoom-out taps are
// adjacent zoomed pixels, whose zoom-in taps start at most one raw pixel apart (the zoom-in
// step is <= 1).  Merged form: base index (<= N-3, so base..base+2 is always in range)
// and three weights.
struct Tap3 {
  float w[3];
  int i;
};
__device__ inline Tap3 merge_taps(const Taps64& t, int N) {
  Tap3 m;
  int lo = 0x7fffffff;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (t.w[k] != 0.0) lo = min(lo, t.i[k]);
  double w0 = 0.0, w1 = 0.0, w2 = 0.0;
  if (lo == 0x7fffffff) {
    lo = max((N >> 1) - 1, 0);  // no tap: point at the image centre, which is always rendered
  } else {
    lo = max(min(lo, N - 3), 0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (t.w[k] != 0.0) {
        const int o = t.i[k] - lo;
        if (o <= 0) w0 += t.w[k];
        else if (o == 1) w1 += t.w[k];
        else w2 += t.w[k];
      }
    }
  }
  m.i = lo;
  m.w[0] = (float)w0; m.w[1] = (float)w1; m.w[2] = (float)w2;
  return m;
}

template <int kCluster, int NS>
__global__ void __launch_bounds__(kThreads, 1) jet2_kernel(const Args a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Fixed2& sm = *reinterpret_cast<Fixed2*>(smem_raw);
  const int N = a.size, pitch = a.pitch, slots = a.slots;
  const int npad = (N + 3) & ~3;
  float* tab_w = reinterpret_cast<float*>(smem_raw + sizeof(Fixed2));  // [3][npad] weights
  int* tab_i = reinterpret_cast<int*>(tab_w + 3 * npad);               // [npad] first tap (absolute)
  float* raw = reinterpret_cast<float*>(tab_i + npad);                 // [H (+1 halo)][pitch]

  const uint32_t s_raw = keep_u32(smem_u32(raw)), s_tabw = keep_u32(smem_u32(tab_w)),
                 s_tabi = keep_u32(smem_u32(tab_i));
  const int tid = keep_s32((int)threadIdx.x), lane = tid & 31, warp = tid >> 5;
  const int rank = (kCluster > 1) ? (int)(blockIdx.x % kCluster) : 0;
  const int cid = blockIdx.x / kCluster, nclusters = gridDim.x / kCluster;
  const int H = N / kCluster;                // rows of this CTA's band (N even when kCluster == 2)
  const int R0 = rank * H;                   // first absolute row of the band
  const int base = (rank == 0) ? 0 : H - 1;  // absolute row stored at raw row 0 (rank 1: halo)
  // remote (peer CTA) shared-memory addresses for st.async
  uint32_t peer_xval = 0, peer_xbar = 0, peer_raw = 0;
  if constexpr (kCluster > 1) {
    peer_xval = cluster_map(smem_u32(&sm.xval[0][0]), (uint32_t)(rank ^ 1));
    peer_xbar = cluster_map(smem_u32(&sm.xbar[0]), (uint32_t)(rank ^ 1));
    peer_raw = cluster_map(smem_u32(raw), (uint32_t)(rank ^ 1));
  }
  int xuse[3] = {0, 0, 0};  // uses so far of each exchange mbarrier (same in both CTAs)
  // swap one 64-bit value with the peer CTA (all threads call; value is CTA-uniform)
  auto exchange = [&](int kind, unsigned long long mine) -> unsigned long long {
    if constexpr (kCluster == 1) {
      return mine;
    } else {
      const int par = xuse[kind] & 1;
      if (tid == 0) {
        st_async_b64(peer_xval + (uint32_t)(kind * 2 + par) * 8u, mine, peer_xbar + (uint32_t)kind * 8u);
        mbar_expect_tx(&sm.xbar[kind], 8);
      }
      mbar_wait(&sm.xbar[kind], (uint32_t)par);
      xuse[kind]++;
      return sm.xval[kind][par];
    }
  };

  const uint32_t pbytes = (uint32_t)slots * RSB_JET_PARAM_STRIDE * 8u;
  const int strips = (N + kStrip - 1) / kStrip;           // 128-column strips (H/V pass)
  const int tcols = (N + kTileCols - 1) / kTileCols;      // 64-column tiles (phase 1)
  const int ntiles = ((H + kTileRows - 1) / kTileRows) * tcols;
  const bool vec = (N & 3) == 0 && (reinterpret_cast<uintptr_t>(a.out) & 15) == 0;
  const bool f64 = (a.flags & RSB_JET_OUT_F64) != 0;
  const bool normalize = (a.flags & RSB_JET_NORMALIZE) != 0;
  const size_t esz = f64 ? 8 : 4;
  // float32 output with contiguous rows: the finished band is staged in shared memory (over
  // the raw image) and leaves through one bulk async copy
  const bool bulk = false && !f64 && vec && pitch == N;  // measured: TMA drains shared memory no faster than STG (~30 B/clk/SM)

  if (tid == 0) {
    mbar_init(&sm.bar[0], 1);
    mbar_init(&sm.bar[1], 1);
    mbar_init(&sm.bar[2], 1);
    mbar_init(&sm.xbar[0], 1);
    mbar_init(&sm.xbar[1], 1);
    mbar_init(&sm.xbar[2], 1);
    fence_mbar_init();
    sm.lmargin = N;
    sm.lbest = 0ull;
    sm.next_tile = 0;
    sm.next_row = 0;
  }
  // tile visiting order: by distance of the tile centre from the image centre, so that the
  // heavy central tiles are dealt first and the dynamic schedule ends on light ones
  for (int t = tid; t < ntiles; t += kWorkers) {
    auto key = [&](int u) {
      const int gg = u / tcols, cc = u - gg * tcols;
      const int dy = 2 * (R0 + gg * kTileRows) + kTileRows - N, dx = 2 * cc * kTileCols + kTileCols - N;
      return (dy * dy + dx * dx) * 256 + u;  // unique
    };
    const int mykey = key(t);
    int pos = 0;
    for (int u = 0; u < ntiles; ++u) pos += key(u) < mykey;
    sm.tile_order[pos] = (unsigned char)t;
  }
  __syncthreads();
  auto prefetch = [&](int image, int pb) {  // one thread: parameter block -> params[pb]
    fence_proxy_async();
    mbar_expect_tx(&sm.bar[pb], pbytes);
    bulk_g2s(sm.params[pb], a.params + (size_t)image * slots * 6, pbytes, &sm.bar[pb]);
  };
  if (tid == 0 && cid < a.n) {
    prefetch(cid, 0);
    if (cid + nclusters < a.n) prefetch(cid + nclusters, 1);
  }
  if (warp == 0 && cid < a.n) {  // prologue: coefficients of the first image
    mbar_wait(&sm.bar[0], 0);
    build_coefs2(sm, 0, sm.params[0], slots, N, R0, H, lane, a.zoom_pad[cid] >= 0);
  }
  // the peer CTA is resident and its mbarriers are initialised before any st.async
  if constexpr (kCluster > 1) cg::this_cluster().sync(); else __syncthreads();
  int zpad_next = (cid < a.n) ? a.zoom_pad[cid] : -1;

  // Two warps start each image with a side job before they join the tile queue: warp 0 builds
  // the NEXT image's coefficients and culling spans, warp 1 turns the PREVIOUS image's arg-max
  // into the float64 peak label.  (Dedicated helper warps were tried: no gain, more spills.)
  constexpr int kCoefWarp = 0, kLabelWarp = 1;
  auto worker_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(kWorkers) : "memory"); };
  auto label_job = [&](int par) {
    const Fixed2::LabelReq q = sm.label[par];
    double peak = 1.0;
    if (normalize)
      peak = peak64(sm.pix64, sm.params[q.pb], slots, N, q.zpad >= 0, q.zpad, q.margin, q.best, lane);
    if (lane == 0) {
      a.margin[q.img] = q.margin;
      a.peak[q.img] = peak;
    }
  };

  int it = 0;
  for (int img = cid; img < a.n; img += nclusters, ++it) {
    const int buf = it & 1;   // coefficient / exchange-slot parity
    const int pb = it % 3;    // parameter buffer
    mbar_wait(&sm.bar[pb], (it / 3) & 1);
    if (tid == 0) bulk_wait_read();  // the previous image's band has left shared memory
    __syncthreads();  // (A) coef/span/nlive of this image visible; raw free for reuse
    if (tid == 0) {  // running reductions of this image (all reads of the previous values are behind A)
      sm.lbest = 0ull;
      sm.lmargin = N;
      sm.ncand = 0;
    }
    __syncwarp();
    {
      if (warp == kCoefWarp) {
        if (img + nclusters < a.n) {
          const int nb = (it + 1) % 3;
          mbar_wait(&sm.bar[nb], ((it + 1) / 3) & 1);
          build_coefs2(sm, buf ^ 1, sm.params[nb], slots, N, R0, H, lane,
                       a.zoom_pad[img + nclusters] >= 0);
        }
        RSB_STAMP_T(12, kCoefWarp * 32);
      } else if (warp == kLabelWarp) {
        // previous image's labels, then its parameter buffer is free for image it+2
        if (rank == 0 && it > 0) label_job(buf ^ 1);
        if (lane == 0 && img + 2 * nclusters < a.n) prefetch(img + 2 * nclusters, (it + 2) % 3);
        RSB_STAMP_T(13, kLabelWarp * 32);
      }
    }
    RSB_STAMP(0);
    const double* P = sm.params[pb];
    const int zpad = zpad_next;
    if (img + nclusters < a.n) zpad_next = a.zoom_pad[img + nclusters];  // latency hidden
    const bool zoom = zpad >= 0;
    const int nlive = sm.nlive[buf];
    const float thr = (float)kEdgeThr, delta = sm.delta[buf];
    const int dlo = sm.dlo[buf];
    const uint32_t s_coef = keep_u32(smem_u32(sm.coef[buf])), s_span = keep_u32(smem_u32(sm.span[buf]));

    // ---------------- phase 1: splat this CTA's band, 8 x 64 tiles ------------------
    int my_margin = N;
    unsigned long long my_best = 0ull;
    float best_v = -1.f;
    const float inv_tcols = 1.f / (float)tcols;
    for (;;) {
      int t = 0;
      if (lane == 0) t = atomicAdd(&sm.next_tile, 1);
      t = __shfl_sync(0xffffffffu, t, 0);
      if (t >= ntiles) break;
      const int ta = sm.tile_order[t];  // central (heavy) tiles first, light corners last
      const int g = __float2int_rz(((float)ta + 0.5f) * inv_tcols), tc = ta - g * tcols;
      const int r0 = R0 + g * kTileRows, tc0 = tc * kTileCols, c0 = tc0 + 2 * lane;
      const uint32_t s_dst = s_raw + (uint32_t)((r0 - base) * pitch + c0) * 4u;
      const int rend = min(R0 + H, r0 + kTileRows);  // rows of this tile that exist
      {
        // every pixel of the tile closer than dlo to the border: never read, never above 0.01
        const int rmid = min(max(N >> 1, r0), rend - 1), cmid = min(max(N >> 1, tc0), min(tc0 + kTileCols, N) - 1);
        const int far = min(min(rmid, N - 1 - rmid), min(cmid, N - 1 - cmid));
        // (two pixels of slack: zero-weight taps may touch the crop's neighbours)
        if (zoom && far < dlo - 2) continue;
      }
      const float r0f = (float)r0, c0f = (float)c0;
      f2 acc2[kTileRows];  // two adjacent pixels per row, packed (FFMA2)
#pragma unroll
      for (int i = 0; i < kTileRows; ++i) acc2[i] = 0ull;
      // cull: lane k tests live slot k (and k+32) against the tile, the warp walks the hits
      for (int k0 = 0; k0 < nlive; k0 += 32) {
        bool hit = false;
        if (k0 + lane < nlive) {
          const int sp = lds_s32(s_span + (uint32_t)((k0 + lane) * kMaxGroups + g) * 4u);
          hit = !((sp >> 16) < tc0 || (sp & 0xffff) > tc0 + kTileCols - 1);
        }
        unsigned hits = __ballot_sync(0xffffffffu, hit);
        while (hits) {
          const int k = k0 + __ffs(hits) - 1;
          hits &= hits - 1;
          const float4 q = lds_v4(s_coef + (uint32_t)k * 48u);        // A B C amp
          const float4 ct = lds_v4(s_coef + (uint32_t)k * 48u + 16u);  // irf fr icf fc
          const float d0 = (c0f - ct.z) - ct.w;
          const f2 d = f2_pack(d0, d0 + 1.f);
          const f2 Q = f2_mul(f2_pack(q.y, q.y), d);
          const f2 Pq = f2_mul(f2_mul(f2_pack(q.x, q.x), d), d);
          const f2 amp2 = f2_pack(q.w, q.w);
          const float dr0 = (r0f - ct.x) - ct.y;
#pragma unroll
          for (int i = 0; i < kTileRows; i += 2) {
            // two rows at a time: dr and C dr^2 as a packed pair, then broadcast per row
            const f2 drp = f2_pack(dr0 + (float)i, dr0 + (float)(i + 1));
            const f2 Cqp = f2_mul(f2_mul(f2_pack(q.z, q.z), drp), drp);
            float dra, drb, Cqa, Cqb;
            f2_unpack(drp, dra, drb);
            f2_unpack(Cqp, Cqa, Cqb);
            float ta, tb;
            f2_unpack(f2_fma(f2_pack(dra, dra), Q, f2_add(Pq, f2_pack(Cqa, Cqa))), ta, tb);
            acc2[i] = f2_fma(amp2, f2_pack(ex2(ta), ex2(tb)), acc2[i]);
            f2_unpack(f2_fma(f2_pack(drb, drb), Q, f2_add(Pq, f2_pack(Cqb, Cqb))), ta, tb);
            acc2[i + 1] = f2_fma(amp2, f2_pack(ex2(ta), ex2(tb)), acc2[i + 1]);
          }
        }
      }
      float acc[kTileRows][2];
#pragma unroll
      for (int i = 0; i < kTileRows; ++i) f2_unpack(acc2[i], acc[i][0], acc[i][1]);
      if (c0 < N) {
        float tile_top = 0.f;
#pragma unroll
        for (int i = 0; i < kTileRows; ++i) {
          if (r0 + i < rend) {
            sts_v2(s_dst + (uint32_t)(i * pitch) * 4u, acc[i][0], acc[i][1]);
            tile_top = fmaxf(tile_top, fmaxf(acc[i][0], acc[i][1]));
          }
        }
        if (zoom) {
          // border distance of this lane's nearest-to-border pixel in the tile
          const int rl = rend - 1;
          const int dmin = min(min(r0, N - 1 - rl), min(c0, N - 1 - min(c0 + 1, N - 1)));
          if (tile_top > thr - delta && dmin < my_margin) {
#pragma unroll
            for (int i = 0; i < kTileRows; ++i) {
              const int r = r0 + i;
              if (r >= rend) break;
              const int db = min(r, N - 1 - r);
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                const int c = c0 + j;
                const int d = min(db, min(c, N - 1 - c));
                if (c < N && d < my_margin && acc[i][j] > thr - delta) {
                  if (acc[i][j] >= thr + delta) {
                    my_margin = d;  // definitely above
                  } else {
                    // within float32 noise of the threshold: float64 decides, later and only
                    // if it can still lower the margin
                    const int slot = atomicAdd(&sm.ncand, 1);
                    if (slot < kMaxCand) sm.cand[slot] = (r << 16) | c;
                    else if (eval64(P, slots, r, c) > kEdgeThr) my_margin = d;
                  }
                }
              }
            }
          }
        } else if (normalize && tile_top >= best_v) {
#pragma unroll
          for (int i = 0; i < kTileRows; ++i) {
            const int r = r0 + i;
            if (r >= rend) break;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              if (c0 + j < N) {
                const unsigned long long key = best_key(acc[i][j], r * N + c0 + j);
                my_best = key > my_best ? key : my_best;
              }
            }
          }
          best_v = __uint_as_float((unsigned)(my_best >> 32));
        }
      }
    }

    RSB_STAMP(1);  // this warp is out of tiles
    int margin = 0;
    char* out = reinterpret_cast<char*>(a.out) + (size_t)img * N * N * esz;
    unsigned long long image_best = 0ull;  // arg-max key of the final image (normalize only)

    if (!zoom) {
      // ------------- plain branch ---------------------------------------------------
      if (normalize) {
        my_best = warp_max_u64(my_best);
        if (lane == 0) atomicMax(&sm.lbest, my_best);
      }
      worker_sync();
      image_best = sm.lbest;
      if (normalize) {
        const unsigned long long theirs = exchange(2, image_best);
        image_best = theirs > image_best ? theirs : image_best;
      }
      RSB_STAMP(2);
      if (tid == 0) sm.next_tile = 0;
      const float inv = normalize ? 1.f / __uint_as_float((unsigned)(image_best >> 32)) : 1.f;
      const int q4 = npad >> 2;
      for (int e = tid; e < H * q4; e += kWorkers) {
        const int rl = e / q4, c0 = (e - rl * q4) << 2;
        const int r = R0 + rl;
        float4 v = lds_v4(s_raw + (uint32_t)((r - base) * pitch + c0) * 4u);
        v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
        if (bulk) *reinterpret_cast<float4*>(raw + (size_t)(r - base) * pitch + c0) = v;
        else store_out4(out + (size_t)r * N * esz, f64, c0, N, vec, v);
      }
    } else {
      // ------------- zoom branch ------------------------------------------------------
      my_margin = warp_min(my_margin);
      if (lane == 0 && my_margin < N) atomicMin(&sm.lmargin, my_margin);
      worker_sync();
      {
        const int ncand = min(sm.ncand, kMaxCand);
        if (ncand > 0) {  // CTA-uniform, rare
          const int mdef = sm.lmargin;
          worker_sync();
          for (int q = warp; q < ncand; q += kWarps) {
            const int rc = sm.cand[q], r = rc >> 16, c = rc & 0xffff;
            const int d = min(min(r, N - 1 - r), min(c, N - 1 - c));
            if (d < mdef) {
              double s = 0.0;
              for (int k0 = 0; k0 < slots; k0 += 32) {
                const double t = (k0 + lane < slots) ? gauss64(P + 6 * (k0 + lane), r, c) : 0.0;
                for (int l = 0; l < 32; ++l) s += __shfl_sync(0xffffffffu, t, l);  // slot order
              }
              if (lane == 0 && s > kEdgeThr) atomicMin(&sm.lmargin, d);
            }
          }
          worker_sync();
        }
      }
      margin = sm.lmargin;
      margin = min(margin, (int)exchange(0, (unsigned long long)(unsigned)margin));
      RSB_STAMP(2);
      if (margin >= N) margin = 0;  // empty mask -> idx 0 (argmax of an all-false array)
      if (tid == 0) {
        sm.next_tile = 0;
        sm.next_row = 0;
        sm.foot_lo = N;
        sm.foot_hi = -1;
      }
      worker_sync();  // foot_* initialised before the atomics below

      // composite tap table (float64 math, float32 weights) + output footprint
      {
        int lo = N, hi = -1;
        for (int d = tid; d < npad; d += kWorkers) {
          Tap3 m;
          m.w[0] = m.w[1] = m.w[2] = 0.f;
          m.i = 0;
          if (d < N) {
            m = merge_taps(composite_taps(d, N, margin, zpad), N);
            if (m.w[0] != 0.f || m.w[1] != 0.f || m.w[2] != 0.f) {
              lo = min(lo, d);
              hi = max(hi, d);
            }
          }
#pragma unroll
          for (int k = 0; k < 3; ++k) tab_w[k * npad + d] = m.w[k];
          tab_i[d] = m.i;
        }
        lo = warp_min(lo);
        hi = -warp_min(-hi);
        if (lane == 0) {
          if (lo < N) atomicMin(&sm.foot_lo, lo);
          if (hi >= 0) atomicMax(&sm.foot_hi, hi);
        }
      }
      worker_sync();  // (D)
      RSB_STAMP(3);
      const int flo = sm.foot_lo, fhi = sm.foot_hi;  // empty footprint: flo = N, fhi = -1


      int dbg_rows_out = 0;
      // H pass, in place: row r of raw becomes row r of T.  Column taps stay in registers,
      // rows are pulled from a shared counter (warp 0 joins late).
      {
        Tap3 ct[NS][4];
        bool act[NS];
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          // lane <-> ADJACENT columns (c = strip + 32 j + lane): the gathers of a warp touch
          // consecutive words (composite stride <= 2), not every 4th
          const int sc0 = s * kStrip;
          act[s] = (s < strips) && sc0 <= fhi && sc0 + kStrip - 1 >= flo;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int cj = sc0 + 32 * j + lane;
            const int c = min(cj, npad - 1);
#pragma unroll
            for (int k = 0; k < 3; ++k)
              ct[s][j].w[k] = (act[s] && cj < N) ? tab_w[k * npad + c] : 0.f;
            ct[s][j].i = act[s] ? tab_i[c] : max((N >> 1) - 1, 0);
          }
        }
        const int rlo = max(R0, margin), rhi = min(R0 + H, N - margin);
        RSB_STAMP(10);
        // rows are dealt statically; inactive strips carry zero weights and index 0, so the
        // gathers need no branches (24 independent loads in flight per row)
        for (int r = rlo + warp; r < rhi; r += kWarps) {
          const uint32_t s_row = s_raw + (uint32_t)((r - base) * pitch) * 4u;
          float v[NS][4][3];
#pragma unroll
          for (int s = 0; s < NS; ++s)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t src = s_row + (uint32_t)ct[s][j].i * 4u;
#ifdef RSB_DBG_NOGATHER
              v[s][j][0] = v[s][j][1] = v[s][j][2] = __uint_as_float(src);
#else
              v[s][j][0] = lds_f32(src);
              v[s][j][1] = lds_f32(src + 4u);
              v[s][j][2] = lds_f32(src + 8u);
#endif
            }
          float res[NS][4];
#pragma unroll
          for (int s = 0; s < NS; ++s)
#pragma unroll
            for (int j = 0; j < 4; ++j)
              res[s][j] = fmaf(ct[s][j].w[2], v[s][j][2],
                               fmaf(ct[s][j].w[1], v[s][j][1], ct[s][j].w[0] * v[s][j][0]));
          __syncwarp();
          ++dbg_rows_out;
#pragma unroll
          for (int s = 0; s < NS; ++s) {
            if (act[s]) {
              const uint32_t s_out = s_row + (uint32_t)(s * kStrip + lane) * 4u;
              if ((s + 1) * kStrip <= N) {  // whole strip inside the image (warp-uniform)
#pragma unroll
                for (int j = 0; j < 4; ++j) sts_f32(s_out + 128u * j, res[s][j]);
              } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (s * kStrip + 32 * j + lane < N) sts_f32(s_out + 128u * j, res[s][j]);
              }
            }
          }
        }
      }
#ifdef RSB_TIMELINE
      if (g_timeline && tid == 64) g_timeline[((size_t)img * 2 + rank) * 16 + 11] = dbg_rows_out;
#endif
      RSB_STAMP(4);
      worker_sync();  // (F) T complete in this band
      RSB_STAMP(5);
      if constexpr (kCluster > 1) {
        // halo: my boundary row of T goes into the peer's halo slot
        const int src = (rank == 0) ? (H - 1) : 1;  // local index of absolute row H-1 / H
        const int dst = (rank == 0) ? 0 : H;        // peer's halo slot
        for (int e = tid; e < (pitch >> 2); e += kWorkers)
          st_async_v4(peer_raw + (uint32_t)(dst * pitch + 4 * e) * 4u,
                      reinterpret_cast<const float4*>(raw + (size_t)src * pitch)[e], peer_xbar + 8u);
        if (tid == 0) mbar_expect_tx(&sm.xbar[1], (uint32_t)pitch * 4u);
        mbar_wait(&sm.xbar[1], (uint32_t)(xuse[1] & 1));
        xuse[1]++;
      }

      RSB_STAMP(6);
      // V pass: results stay in registers (<= 8 rows x NS strips of 4 columns per lane)
      const int vr0 = max(R0, flo), vr1 = min(R0 + H - 1, fhi);  // inclusive footprint rows
      float4 res[kRowsPerWarp][NS];
      my_best = 0ull;
      best_v = -1.f;
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) {
        const int r = R0 + warp + i * kWarps;
#pragma unroll
        for (int s = 0; s < NS; ++s) res[i][s] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < vr0 || r > vr1) continue;  // warp-uniform
        const float w0 = lds_f32(s_tabw + (uint32_t)r * 4u), w1 = lds_f32(s_tabw + (uint32_t)(npad + r) * 4u),
                    w2 = lds_f32(s_tabw + (uint32_t)(2 * npad + r) * 4u);
        // only zero-weight taps can fall outside this band (+ halo): clamp them
        const int l0 = lds_s32(s_tabi + (uint32_t)r * 4u) - base;
        const uint32_t p0 = s_raw + (uint32_t)(min(max(l0, 0), H) * pitch) * 4u;
        const uint32_t p1 = s_raw + (uint32_t)(min(max(l0 + 1, 0), H) * pitch) * 4u;
        const uint32_t p2 = s_raw + (uint32_t)(min(max(l0 + 2, 0), H) * pitch) * 4u;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          const int sc0 = s * kStrip, c0 = sc0 + 4 * lane;
          if (!(s < strips && sc0 <= fhi && sc0 + kStrip - 1 >= flo) || c0 >= N) continue;
          const float4 t0 = lds_v4(p0 + (uint32_t)c0 * 4u);
          const float4 t1 = lds_v4(p1 + (uint32_t)c0 * 4u);
          const float4 t2 = lds_v4(p2 + (uint32_t)c0 * 4u);
          float4 v;
          {
            const f2 W0 = f2_pack(w0, w0), W1 = f2_pack(w1, w1), W2 = f2_pack(w2, w2);
            f2_unpack(f2_fma(W2, f2_pack(t2.x, t2.y), f2_fma(W1, f2_pack(t1.x, t1.y), f2_mul(W0, f2_pack(t0.x, t0.y)))), v.x, v.y);
            f2_unpack(f2_fma(W2, f2_pack(t2.z, t2.w), f2_fma(W1, f2_pack(t1.z, t1.w), f2_mul(W0, f2_pack(t0.z, t0.w)))), v.z, v.w);
          }
          res[i][s] = v;
          if (normalize) {
            const float top = fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w));
            if (top >= best_v) {
              const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                if (c0 + j < N) {
                  const unsigned long long key = best_key(vv[j], r * N + c0 + j);
                  my_best = key > my_best ? key : my_best;
                }
              }
              best_v = __uint_as_float((unsigned)(my_best >> 32));
            }
          }
        }
      }
      RSB_STAMP(7);
      float inv = 1.f;
      if (normalize) {
        my_best = warp_max_u64(my_best);
        if (lane == 0) atomicMax(&sm.lbest, my_best);
        worker_sync();  // (G) this band's arg-max complete
        image_best = sm.lbest;
        const unsigned long long theirs = exchange(2, image_best);
        image_best = theirs > image_best ? theirs : image_best;
        inv = 1.f / __uint_as_float((unsigned)(image_best >> 32));
      }
      RSB_STAMP(8);
      if (bulk && !normalize) worker_sync();  // every V-pass read of raw is done
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) {
        const int r = R0 + warp + i * kWarps;
        if (r >= R0 + H) continue;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          const int c0 = s * kStrip + 4 * lane;
          if (s >= strips || c0 >= N) continue;
          float4 v = res[i][s];
          v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
          if (bulk) *reinterpret_cast<float4*>(raw + (size_t)(r - base) * pitch + c0) = v;
          else store_out4(out + (size_t)r * N * esz, f64, c0, N, vec, v);
        }
      }
    }
    if (bulk) {
      fence_proxy_async();  // my shared-memory writes -> visible to the async (TMA) proxy
      worker_sync();
      if (tid == 0)
        bulk_s2g(out + (size_t)R0 * N * 4, raw + (size_t)(R0 - base) * pitch, (uint32_t)(H * N * 4));
    }

    RSB_STAMP(9);
    // this image's labels are produced by the label warp during the next image
    if (tid == 0) {
      Fixed2::LabelReq q;
      q.best = image_best; q.img = img; q.pb = pb; q.margin = margin; q.zpad = zpad;
      sm.label[buf] = q;
    }
  }
  if (tid == 0) bulk_wait_read();
  __syncthreads();
  if (rank == 0 && warp == kLabelWarp && it > 0) label_job((it - 1) & 1);
  if constexpr (kCluster > 1) cg::this_cluster().sync();  // no CTA exits while its smem is in use
}

// ------------------------------------------------------------------ host side
template <int kCluster, int NS>
static int launch_v2(const Args& a, size_t smem_bytes, cudaStream_t st) {
  auto k = jet2_kernel<kCluster, NS>;
  RSB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  const int sms = sm_count();
  // CTAs are limited to one per SM by registers (512 threads x 128); small images leave
  // shared memory for more, but the register file does not
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
  return sizeof(v2::Fixed2) + (size_t)npad * 16 + (size_t)rows * npad * 4;
}

bool v2_supported(int size, int* cluster, size_t* smem_bytes) {
  int dev = 0, max_optin = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess)
    return false;
  if (size <= v2::kMaxBand && v2_bytes(size, 1) <= (size_t)max_optin) {
    *cluster = 1;
    *smem_bytes = v2_bytes(size, 1);
    return true;
  }
  if ((size & 1) == 0 && size / 2 <= v2::kMaxBand && v2_bytes(size, 2) <= (size_t)max_optin) {
    *cluster = 2;
    *smem_bytes = v2_bytes(size, 2);
    return true;
  }
  return false;
}

int v2_launch(const Args& a, int cluster, size_t smem_bytes, cudaStream_t st) {
  const int strips = (a.size + kStrip - 1) / kStrip;  // <= 2 for size <= 256
  if (cluster == 1)
    return strips <= 1 ? v2::launch_v2<1, 1>(a, smem_bytes, st) : v2::launch_v2<1, 2>(a, smem_bytes, st);
  return strips <= 1 ? v2::launch_v2<2, 1>(a, smem_bytes, st) : v2::launch_v2<2, 2>(a, smem_bytes, st);
}

}  // namespace jet
}  // namespace rsb

#ifdef RSB_TIMELINE
extern "C" int rsb_debug_set_timeline(long long* dev_ptr) {
  RSB_CUDA(cudaMemcpyToSymbol(rsb::jet::v2::g_timeline, &dev_ptr, sizeof(dev_ptr)));
  return 0;
}
#endif
