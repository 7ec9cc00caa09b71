// On-chip jet renderer: the raw image never leaves shared memory.
// (reference: src/radiosim/jets/jet.py:144-192, utils/gauss.py:50-97, utils/utils.py:90-177)
//
// One image per thread-block CLUSTER.  kCluster == 1: the whole raw image sits in one
// CTA's shared memory (size <= 128).  kCluster == 2: the image is split into two row
// bands, one per CTA of a 2-CTA cluster (size 256: 129 x 256 floats = 132 KB each);
// the composite zoom operator is centre preserving, so each CTA resamples its own band
// and the only data that crosses CTAs is ONE halo row, the margin and the arg-max,
// written into the peer's shared memory with st.async (completion on the peer's mbarrier).
//
// Per image (DESIGN.md 4.1 has the long version):
//   ahead   : 1-D bulk async copy of the parameter block (triple buffered).  Warp 0 builds,
//             one image AHEAD and off the critical path, the float32 conic coefficients,
//             which slots touch which 8 x 64 tile (box-vs-ellipse test), the frame bound
//             dlo, and the per-warp work lists (tiles sorted by cost, snake deal).
//   phase 1 : every warp walks its own list; 16 pixels per lane, rows in blocks of four
//             from two exponentials per column, next record and next coefficients
//             requested one step ahead.
//   zoom    : margin -> tap table -> H pass in place (column taps in registers) -> halo
//             row exchange -> ONE V pass whose results stay in registers -> arg-max ->
//             scaled 128-bit stores, half of the band through a staged bulk copy that
//             drains under the next image.  The float64 peak for the labels is evaluated
//             by warp 1 (first CTA) while the others already render the next image.
#include <cooperative_groups.h>

#include "jet_common.cuh"

namespace cg = cooperative_groups;

namespace rsb {
namespace jet {
namespace v2 {

#ifdef RSB_TIMELINE
__device__ long long* g_timeline = nullptr;  // [image][rank][16] clock64 stamps (debug builds)
#define RSB_STAMP_T(k, t) do { if (g_timeline && tid == (t)) g_timeline[((size_t)img * 2 + rank) * 96 + (k)] = clock64(); } while (0)
#define RSB_STAMP(k) do { if (g_timeline && tid == 64) g_timeline[((size_t)img * 2 + rank) * 96 + (k)] = clock64(); } while (0)
#else
#define RSB_STAMP(k) do {} while (0)
#define RSB_STAMP_T(k, t) do {} while (0)
#endif
#ifdef RSB_DBG_NOEX2  // experiment: phase 1 without the SFU (wrong images)
#define P1_EX2(x) (x)
#else
#define P1_EX2(x) ex2(x)
#endif

constexpr int kWorkers = 512;             // threads that render
constexpr int kWarps = kWorkers / 32;
constexpr int kThreads = kWorkers;
constexpr int kTileRows = 8;
constexpr int kTileCols = 64;   // 2 columns per lane
constexpr int kMaxBand = 128;   // rows per CTA band (V-pass results live in registers)
constexpr int kBigMax = 224;    // largest image one CTA holds whole (V pass twice; 4 x 32 tiles of 8 x 64)
constexpr int kRowsPerWarp = kMaxBand / kWarps;
constexpr int kMaxGroups = kMaxBand / kTileRows;
constexpr int kMaxCand = 64;
constexpr int kMaxTiles = kMaxGroups * 4;  // 64-column tiles of a band (size <= 256)
constexpr int kMaxRec = 8;                 // tiles per warp and image (work list length)
constexpr unsigned kRecEnd = 0xffffu;      // list terminator in rec.y
constexpr int kListCap = 512;              // images of one launch a cluster can list (else: plain striding)

// float32 coefficients of one live component as the tile loop reads them (warp-uniform):
//   g(r,c) = 2^(la + A dc^2 + B dc dr + C dr^2),  dc = c - (icf + fc), dr = r - (irf + fr)
// Rows are rendered in blocks of four from two exponentials per column,
//   E(i) = E0 U^i 2^(C (i^2 - i)),  E0 = g(r_b, c),  U = 2^(B dc + 2 C dr_b + C),
// so k2 = 2^(2C) and k3 = 2^(6C) ride along.  Components this does not suit (amplitude <= 0,
// sigma below ~0.6 pixel, non-finite numbers) carry k2 < 0 and la = the amplitude itself:
// they take one exponential per pixel.
struct __align__(16) Coef2 {
  float A, B, C, la;
  float irf, fr, icf, fc;  // centre split into integer part and fraction
  float k2, k3;
  int r0, r1, c0, c1;      // inclusive bounding box of the > tau footprint (builder only)
  int pad[2];
};
static_assert(sizeof(Coef2) == 64, "Coef2 layout");

constexpr int kSlots2 = 32;  // the launcher keeps slots <= 32 on this path (per-tile slot sets are 32-bit masks)

struct __align__(16) Fixed2 {
  double params[3][kSlots2 * RSB_JET_PARAM_STRIDE];  // triple buffered (bulk async copies)
  Coef2 coef[2][kSlots2];                              // double buffered, built one image ahead
  int hist[32];                     // build_coefs2 scratch: tiles per cost
  double pix64[22 * kSlots2 + 8];   // peak64 scratch: 16 pixel x slot terms, 6 per-slot constants, tap coordinates
  uint64_t bar[3];
  double scale[2];                  // composite operator: (N + 2 pad) / N and (N - 2 margin) / N of this image
  unsigned long long xval[3][2];    // peer's margin / (unused) / arg-max key, [kind][use parity]
  uint64_t xbar[3];                 // mbarriers the peer's st.async completes: margin, halo, best
  unsigned long long lbest;         // this CTA's running arg-max key
  int lmargin;                      // this CTA's running margin
  int nlive[2];
  int next_row;
  int foot_lo, foot_hi;             // output coordinates with a non-zero tap
  float delta[2];
  int dlo[2];                       // no pixel closer than this to the border exceeds 0.01
  struct LabelReq { unsigned long long best; int img, pb, margin, zpad, n, pad; } label[2];  // [image parity]
  float geo[kSlots2 * 16];          // build_coefs2 scratch (SlotGeo per live slot)
  int ncand;                        // borderline pixels awaiting a float64 verdict
  int cand[kMaxCand];               // r << 16 | c
  uint2 rec[2][kWarps][kMaxRec];    // per-warp work lists: {slots that touch the tile, row group << 8 | column tile}
  int img_list[kListCap];           // this cluster's images: the zoom ones, then the plain ones
  int nimg;                         // ... how many
  int wz[kWarps], wp[kWarps];       // (building it: zoom / plain images per warp's range)
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

// Index of the lowest set bit without FLO/POPC (those are SFU-queue instructions and would
// wait behind the exponentials): isolate the bit, convert to float on the FMA pipe, read the
// exponent.  Negative for x == 0.
__device__ __forceinline__ int lowest_bit(unsigned x) {
  return (int)(__float_as_uint(__uint2float_rn(x & (0u - x))) >> 23) - 127;
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

// ---- one warp: parameter block -> compacted coefficients + per-warp work lists + dlo -----
// Runs while the other warps render the previous image, so it is written for latency: float32
// throughout (only the centre split needs the float64 parameters), every lane busy, and as
// few SFU-queue round trips as possible (divisions, square roots, float->int conversions,
// FLO and POPC all queue behind the renderers' exponentials).
// Ellipse of the > tau footprint: z^T M z <= T^2 with z = (dr, dc).  For a row range the
// covered columns are [min cl, max cr], cl/cr = centre line -/+ half width; cr is concave in
// dr, so its maximum over the range is the ellipse's rightmost point if that row lies in the
// range, else the larger end value (mirror image for cl).
struct __align__(16) SlotGeo {  // per live slot, shared between the lanes through shared memory
  float Mrr, Mcc, Mrc, T2;
  float ec, x, y, det;
  float inv_cc, inv_rr, pad0;
  int degenerate;
  float ka, kx, ky, pad;  // frame bound: amplitude and 1-D decay rates (log2 units)
};
static_assert(sizeof(SlotGeo) == 64, "SlotGeo layout");

// round to nearest on the FMA pipe (|x| < 2^22); the callers widen by one pixel instead of
// asking for floor / ceil
__device__ __forceinline__ int round_fma(float x) {
  return __float_as_int(fminf(fmaxf(x, -4.0e6f), 4.0e6f) + 12582912.f) - 0x4B400000;
}

#ifdef RSB_TIMELINE
#define RSB_BSTAMP(k) do { if (dbg_tl && lane == 0) dbg_tl[k] = clock64(); } while (0)
#else
#define RSB_BSTAMP(k) do {} while (0)
#endif
__device__ inline void build_coefs2(Fixed2& sm, int cb, const double* P, int slots, int N, int R0,
                                    int H, int lane, bool zoom, int first_worker,
                                    long long* dbg_tl = nullptr) {
  RSB_BSTAMP(0);
  SlotGeo* geo = reinterpret_cast<SlotGeo*>(sm.geo);
  float amax = 0.f;
  for (int k = lane; k < slots; k += 32) amax = fmaxf(amax, fabsf((float)P[k * 6]));
  amax = warp_max(amax);
  const float tau = 2e-8f * amax;  // <= 19 x 2e-8 of the peak is dropped in total
  const int ngroups = (H + kTileRows - 1) / kTileRows;
  const int tcols = (N + kTileCols - 1) / kTileCols, ntiles = ngroups * tcols;
  sm.hist[lane] = 0;
  int base = 0;
  bool certified = false, all_nonneg = true;
  for (int k0 = 0; k0 < slots; k0 += 32) {
    const int k = k0 + lane;
    bool live = false;
    Coef2 c;
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
        {
          const bool blocks = amp > 0.f && c.C >= -2.f && isfinite(c.A) && isfinite(c.B) && isfinite(c.C);
          c.la = blocks ? log2f(amp) : amp;
          c.k2 = blocks ? exp2f(2.f * c.C) : -1.f;
          c.k3 = blocks ? exp2f(6.f * c.C) : -1.f;
          c.pad[0] = c.pad[1] = 0;
        }
        const double yr = rint(yd), xr = rint(xd);
        c.irf = (float)yr;
        c.fr = (float)(yd - yr);
        c.icf = (float)xr;
        c.fc = (float)(xd - xr);
        g.x = (float)xd;
        g.y = (float)yd;
        g.T2 = 2.f * logf(fabsf(amp) / tau);
        g.det = g.Mrr * g.Mcc - g.Mrc * g.Mrc;
        const float inv_det = 1.f / g.det, inv_rr = 1.f / g.Mrr;
        g.inv_cc = 1.f / g.Mcc;
        const float er = sqrtf(g.T2 * g.Mcc * inv_det);
        g.ec = sqrtf(g.T2 * g.Mrr * inv_det);
        g.degenerate = !(isfinite(er) && isfinite(g.ec));
        g.inv_rr = inv_rr;
        g.pad0 = 0.f;
        const float lo_r = fmaxf(floorf(g.y - er) - 1.f, 0.f), hi_r = fminf(ceilf(g.y + er) + 1.f, N - 1.f);
        const float lo_c = fmaxf(floorf(g.x - g.ec) - 1.f, 0.f), hi_c = fminf(ceilf(g.x + g.ec) + 1.f, N - 1.f);
        if (g.degenerate) {
          live = true;
          c.r0 = 0; c.r1 = N - 1; c.c0 = 0; c.c1 = N - 1;
        } else if (lo_r <= hi_r && lo_c <= hi_c) {
          live = true;
          c.r0 = (int)lo_r; c.r1 = (int)hi_r; c.c0 = (int)lo_c; c.c1 = (int)hi_c;
        }
        g.ka = fabsf(amp) * 1.0001f;
        g.kx = (g.det * inv_rr) * 0.9999f * hl;
        g.ky = (g.det * g.inv_cc) * 0.9999f * hl;
        g.pad = 0.f;
        // the pixel nearest to the centre is at most half a pixel away in each coordinate:
        // this slot alone lifts it above the threshold -> the > 0.01 mask is not empty
        const float qmax = 0.25f * (g.Mrr + g.Mcc + 2.f * fabsf(g.Mrc));
        if (amp * __expf(-0.5f * qmax) > 0.0101f && g.x >= 0.f && g.x <= N - 1.f && g.y >= 0.f &&
            g.y <= N - 1.f)
          certified = true;
      }
      if (p[0] < 0.0) all_nonneg = false;
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
  RSB_BSTAMP(1);
  RSB_BSTAMP(2);
  // Zoom images only use the crop [margin, N-margin)^2 of the raw image, and margin >= dlo,
  // where dlo is a border distance d such that an UPPER BOUND of the image over the frame
  // {border distance <= d - 1} stays below the 0.01 threshold.  Bound: sum over slots of the
  // slot's maximum over the frame = amp if its centre lies in the frame, else amp * exp(-q/2)
  // with q the smallest Mahalanobis distance to the four frame lines.  The bound grows with d:
  // 32 frames are tried at once, one per lane (dlo only feeds the tile skip, a multiple of
  // `step` is as good as the exact value).
  int dlo = 0;
  all_nonneg = __all_sync(0xffffffffu, all_nonneg);
  if (zoom) {
    certified = __any_sync(0xffffffffu, certified) && all_nonneg;
    // without the certificate an empty mask (margin := 0, whole image used) cannot be ruled out
    if (certified) {
      const int half = (N - 1) / 2, step = (half + 31) / 32;
      const int d = step * lane + step - 1;
      const float lo = (float)d, hi = (float)(N - 1 - d);
      float u = 0.f;
#pragma unroll 4
      for (int pos = 0; pos < base; ++pos) {  // branch-free: the exponentials of a batch overlap
        const float4 gx = *reinterpret_cast<const float4*>(&geo[pos].ec);   // ec x y det
        const float4 gk = *reinterpret_cast<const float4*>(&geo[pos].ka);   // ka kx ky -
        const bool in_frame = gx.y <= lo || gx.y >= hi || gx.z <= lo || gx.z >= hi;
        const float dc0 = lo - gx.y, dc1 = hi - gx.y, dr0 = lo - gx.z, dr1 = hi - gx.z;
        const float q = fminf(fminf(dc0 * dc0, dc1 * dc1) * gk.y, fminf(dr0 * dr0, dr1 * dr1) * gk.z);
        u += gk.x * ex2(in_frame ? 0.f : -q);
      }
      const unsigned okm = __ballot_sync(0xffffffffu, d <= half && u <= 0.00999f);
      const unsigned bad = ~okm;
      const int lead = bad ? lowest_bit(bad) : 32;  // leading frames that hold nothing above the threshold
      dlo = (lead == 32) ? 0 : step * lead;         // (all 32 below: cannot happen with a certificate; stay safe)
    }
  }
  if (lane == 0) {
    sm.nlive[cb] = base;
    // float32 noise of a rendered pixel near the 0.01 threshold.  All amplitudes >= 0: every
    // term's error is relative to the term (<= 3e-5: SFU 2^-22 per factor, argument rounding
    // with cancellation, accumulation), so the sum's is relative to the pixel value; plus
    // the tails dropped by the footprint cut (slots x tau).  Otherwise: relative to the
    // largest amplitude.
    sm.delta[cb] = all_nonneg ? 3e-7f + 4e-7f * amax : 6e-6f * fmaxf(1.f, amax);
    sm.dlo[cb] = dlo;
  }
  RSB_BSTAMP(3);
  // ---- static work lists -------------------------------------------------------------
  // Every tile's set of overlapping slots is known here, one image ahead of its use, so the
  // tiles are dealt to the warps now: sorted by cost (slots hit), heaviest first, in snake
  // order.  The renderers then read their list with plain loads -- no queue, no atomics, no
  // shuffles on their critical path.  Warps 0 and 1 (side jobs) receive none.
  {
    uint2* rec = &sm.rec[cb][0][0];
    for (int e = lane; e < kWarps * kMaxRec; e += 32) rec[e] = make_uint2(0u, kRecEnd);
    __syncwarp();
    unsigned hits[2] = {0u, 0u};
    int cost[2] = {0, 0}, tag[2] = {0, 0}, among[2] = {0, 0};
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int u = lane + 32 * h;
      if (u < ntiles) {
        const unsigned inv_tc = (65536u + (unsigned)tcols - 1u) / (unsigned)tcols;
        const int g = (int)(((unsigned)u * inv_tc) >> 16), tc = u - g * tcols;
        const int r0 = R0 + g * kTileRows, tc0 = tc * kTileCols;
        const int rend = min(R0 + H, r0 + kTileRows);
        // every pixel of the tile closer than dlo to the border: never read, never above 0.01
        // (two pixels of slack: zero-weight taps may touch the crop's neighbours)
        const int rmid = min(max(N >> 1, r0), rend - 1), cmid = min(max(N >> 1, tc0), min(tc0 + kTileCols, N) - 1);
        const int far = min(min(rmid, N - 1 - rmid), min(cmid, N - 1 - cmid));
        if (!(zoom && far < dlo - 2)) {
          // a slot touches the tile when its > tau ellipse  M[dr,dc] <= T^2  meets the tile's
          // rectangle (1.5 pixels of slack): the form's minimum over a box that does not hold
          // the centre lies on one of the four edges.  No square roots, no divisions.
          const float ra = (float)r0 - 1.5f, rb = (float)(rend - 1) + 1.5f;
          const float ca = (float)tc0 - 1.5f, cb_ = (float)(min(tc0 + kTileCols, N) - 1) + 1.5f;
          unsigned m = 0u;
#pragma unroll 4
          for (int pos = 0; pos < base; ++pos) {  // branch-free: the loads of a batch overlap
            const int2 bbr = *reinterpret_cast<const int2*>(&sm.coef[cb][pos].r0);  // r0 r1
            const int2 bbc = *reinterpret_cast<const int2*>(&sm.coef[cb][pos].c0);  // c0 c1
            const int4 bb = make_int4(bbr.x, bbr.y, bbc.x, bbc.y);
            const float4 gm = *reinterpret_cast<const float4*>(&geo[pos].Mrr);     // Mrr Mcc Mrc T2
            const float4 gx = *reinterpret_cast<const float4*>(&geo[pos].ec);      // ec x y det
            const float4 gi = *reinterpret_cast<const float4*>(&geo[pos].inv_cc);  // inv_cc inv_rr - -
            const bool inbox = !(bb.y < r0 || bb.x >= rend || bb.w < tc0 || bb.z >= tc0 + kTileCols);
            const float dra = ra - gx.z, drb = rb - gx.z, dca = ca - gx.y, dcb = cb_ - gx.y;
            // only the edges that face the centre can hold the minimum
            const bool use_c = dca > 0.f || dcb < 0.f, use_r = dra > 0.f || drb < 0.f;
            const float e = dca > 0.f ? dca : dcb, f = dra > 0.f ? dra : drb;
            const float tc_ = fminf(fmaxf(-gm.z * e * gi.y, dra), drb);  // column edge dc = e, dr free
            const float qc = (gm.x * tc_ + 2.f * gm.z * e) * tc_ + gm.y * e * e;
            const float tr_ = fminf(fmaxf(-gm.z * f * gi.x, dca), dcb);  // row edge dr = f, dc free
            const float qr = (gm.y * tr_ + 2.f * gm.z * f) * tr_ + gm.x * f * f;
            const float big = 3.0e38f;
            float qmin = fminf(use_c ? qc : big, use_r ? qr : big);
            if (!use_c && !use_r) qmin = 0.f;  // the centre lies inside the rectangle
            // (NaN from a degenerate slot counts as touching)
            if (inbox && !(qmin > gm.w)) m |= 1u << pos;
          }
          hits[h] = m;
          cost[h] = 1 + min(__popc(hits[h]), 30);
          tag[h] = (g << 8) | tc;
          among[h] = atomicAdd(&sm.hist[cost[h]], 1);  // position among the tiles of equal cost
        }
      }
    }
    __syncwarp();
    // rank = tiles of higher cost + position among equals
    int above[2] = {0, 0};
    for (int c = 31; c >= 1; --c) {
      const int n = sm.hist[c];
#pragma unroll
      for (int h = 0; h < 2; ++h)
        if (c > cost[h]) above[h] += n;
    }
    // Snake deal over the warps without a side job: first_worker .. kWarps - 1.
    const int seats = kWarps - first_worker;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (cost[h] == 0) continue;
      const int i = above[h] + among[h];
      const int ps = (i >= seats) + (i >= 2 * seats) + (i >= 3 * seats) + (i >= 4 * seats) + (i >= 5 * seats);
      const int j = i - ps * seats;
      const int w = first_worker + ((ps & 1) ? j : seats - 1 - j), e = ps;
      sm.rec[cb][w][e] = make_uint2(hits[h], (unsigned)tag[h]);
    }
  }
  RSB_BSTAMP(4);
}

// TPL = tiles per lane the work-list part can hold: 2 (bands of up to 128 rows x 256 columns) or 4 (one CTA
// with the whole image, up to 224 rows)
template <int TPL>
__device__ inline void build_coefs2_big(Fixed2& sm, int cb, const double* P, int slots, int N, int R0,
                                    int H, int lane, bool zoom, int first_worker,
                                    long long* dbg_tl = nullptr) {
  RSB_BSTAMP(0);
  SlotGeo* geo = reinterpret_cast<SlotGeo*>(sm.geo);
  float amax = 0.f;
  for (int k = lane; k < slots; k += 32) amax = fmaxf(amax, fabsf((float)P[k * 6]));
  amax = warp_max(amax);
  const float tau = 2e-8f * amax;  // <= 19 x 2e-8 of the peak is dropped in total
  const int ngroups = (H + kTileRows - 1) / kTileRows;
  const int tcols = (N + kTileCols - 1) / kTileCols, ntiles = ngroups * tcols;
  sm.hist[lane] = 0;
  int base = 0;
  bool certified = false, all_nonneg = true;
  for (int k0 = 0; k0 < slots; k0 += 32) {
    const int k = k0 + lane;
    bool live = false;
    Coef2 c;
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
        {
          const bool blocks = amp > 0.f && c.C >= -2.f && isfinite(c.A) && isfinite(c.B) && isfinite(c.C);
          c.la = blocks ? log2f(amp) : amp;
          c.k2 = blocks ? exp2f(2.f * c.C) : -1.f;
          c.k3 = blocks ? exp2f(6.f * c.C) : -1.f;
          c.pad[0] = c.pad[1] = 0;
        }
        const double yr = rint(yd), xr = rint(xd);
        c.irf = (float)yr;
        c.fr = (float)(yd - yr);
        c.icf = (float)xr;
        c.fc = (float)(xd - xr);
        g.x = (float)xd;
        g.y = (float)yd;
        g.T2 = 2.f * logf(fabsf(amp) / tau);
        g.det = g.Mrr * g.Mcc - g.Mrc * g.Mrc;
        const float inv_det = 1.f / g.det, inv_rr = 1.f / g.Mrr;
        g.inv_cc = 1.f / g.Mcc;
        const float er = sqrtf(g.T2 * g.Mcc * inv_det);
        g.ec = sqrtf(g.T2 * g.Mrr * inv_det);
        g.degenerate = !(isfinite(er) && isfinite(g.ec));
        g.inv_rr = inv_rr;
        g.pad0 = 0.f;
        const float lo_r = fmaxf(floorf(g.y - er) - 1.f, 0.f), hi_r = fminf(ceilf(g.y + er) + 1.f, N - 1.f);
        const float lo_c = fmaxf(floorf(g.x - g.ec) - 1.f, 0.f), hi_c = fminf(ceilf(g.x + g.ec) + 1.f, N - 1.f);
        if (g.degenerate) {
          live = true;
          c.r0 = 0; c.r1 = N - 1; c.c0 = 0; c.c1 = N - 1;
        } else if (lo_r <= hi_r && lo_c <= hi_c) {
          live = true;
          c.r0 = (int)lo_r; c.r1 = (int)hi_r; c.c0 = (int)lo_c; c.c1 = (int)hi_c;
        }
        g.ka = fabsf(amp) * 1.0001f;
        g.kx = (g.det * inv_rr) * 0.9999f * hl;
        g.ky = (g.det * g.inv_cc) * 0.9999f * hl;
        g.pad = 0.f;
        // the pixel nearest to the centre is at most half a pixel away in each coordinate:
        // this slot alone lifts it above the threshold -> the > 0.01 mask is not empty
        const float qmax = 0.25f * (g.Mrr + g.Mcc + 2.f * fabsf(g.Mrc));
        if (amp * __expf(-0.5f * qmax) > 0.0101f && g.x >= 0.f && g.x <= N - 1.f && g.y >= 0.f &&
            g.y <= N - 1.f)
          certified = true;
      }
      if (p[0] < 0.0) all_nonneg = false;
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
  RSB_BSTAMP(1);
  RSB_BSTAMP(2);
  // Zoom images only use the crop [margin, N-margin)^2 of the raw image, and margin >= dlo,
  // where dlo is a border distance d such that an UPPER BOUND of the image over the frame
  // {border distance <= d - 1} stays below the 0.01 threshold.  Bound: sum over slots of the
  // slot's maximum over the frame = amp if its centre lies in the frame, else amp * exp(-q/2)
  // with q the smallest Mahalanobis distance to the four frame lines.  The bound grows with d:
  // 32 frames are tried at once, one per lane (dlo only feeds the tile skip, a multiple of
  // `step` is as good as the exact value).
  int dlo = 0;
  all_nonneg = __all_sync(0xffffffffu, all_nonneg);
  if (zoom) {
    certified = __any_sync(0xffffffffu, certified) && all_nonneg;
    // without the certificate an empty mask (margin := 0, whole image used) cannot be ruled out
    if (certified) {
      const int half = (N - 1) / 2, step = (half + 31) / 32;
      const int d = step * lane + step - 1;
      const float lo = (float)d, hi = (float)(N - 1 - d);
      float u = 0.f;
#pragma unroll 4
      for (int pos = 0; pos < base; ++pos) {  // branch-free: the exponentials of a batch overlap
        const float4 gx = *reinterpret_cast<const float4*>(&geo[pos].ec);   // ec x y det
        const float4 gk = *reinterpret_cast<const float4*>(&geo[pos].ka);   // ka kx ky -
        const bool in_frame = gx.y <= lo || gx.y >= hi || gx.z <= lo || gx.z >= hi;
        const float dc0 = lo - gx.y, dc1 = hi - gx.y, dr0 = lo - gx.z, dr1 = hi - gx.z;
        const float q = fminf(fminf(dc0 * dc0, dc1 * dc1) * gk.y, fminf(dr0 * dr0, dr1 * dr1) * gk.z);
        u += gk.x * ex2(in_frame ? 0.f : -q);
      }
      const unsigned okm = __ballot_sync(0xffffffffu, d <= half && u <= 0.00999f);
      const unsigned bad = ~okm;
      const int lead = bad ? lowest_bit(bad) : 32;  // leading frames that hold nothing above the threshold
      dlo = (lead == 32) ? 0 : step * lead;         // (all 32 below: cannot happen with a certificate; stay safe)
    }
  }
  if (lane == 0) {
    sm.nlive[cb] = base;
    // float32 noise of a rendered pixel near the 0.01 threshold.  All amplitudes >= 0: every
    // term's error is relative to the term (<= 3e-5: SFU 2^-22 per factor, argument rounding
    // with cancellation, accumulation), so the sum's is relative to the pixel value; plus
    // the tails dropped by the footprint cut (slots x tau).  Otherwise: relative to the
    // largest amplitude.
    sm.delta[cb] = all_nonneg ? 3e-7f + 4e-7f * amax : 6e-6f * fmaxf(1.f, amax);
    sm.dlo[cb] = dlo;
  }
  RSB_BSTAMP(3);
  // ---- static work lists -------------------------------------------------------------
  // Every tile's set of overlapping slots is known here, one image ahead of its use, so the
  // tiles are dealt to the warps now: sorted by cost (slots hit), heaviest first, in snake
  // order.  The renderers then read their list with plain loads -- no queue, no atomics, no
  // shuffles on their critical path.  Warps 0 and 1 (side jobs) receive none.
  {
    uint2* rec = &sm.rec[cb][0][0];
    for (int e = lane; e < kWarps * kMaxRec; e += 32) rec[e] = make_uint2(0u, kRecEnd);
    __syncwarp();
    unsigned hits[TPL];
    int cost[TPL], tag[TPL], among[TPL];
#pragma unroll
    for (int h = 0; h < TPL; ++h) { hits[h] = 0u; cost[h] = tag[h] = among[h] = 0; }
#pragma unroll
    for (int h = 0; h < TPL; ++h) {
      const int u = lane + 32 * h;
      if (u < ntiles) {
        const unsigned inv_tc = (65536u + (unsigned)tcols - 1u) / (unsigned)tcols;
        const int g = (int)(((unsigned)u * inv_tc) >> 16), tc = u - g * tcols;
        const int r0 = R0 + g * kTileRows, tc0 = tc * kTileCols;
        const int rend = min(R0 + H, r0 + kTileRows);
        // every pixel of the tile closer than dlo to the border: never read, never above 0.01
        // (two pixels of slack: zero-weight taps may touch the crop's neighbours)
        const int rmid = min(max(N >> 1, r0), rend - 1), cmid = min(max(N >> 1, tc0), min(tc0 + kTileCols, N) - 1);
        const int far = min(min(rmid, N - 1 - rmid), min(cmid, N - 1 - cmid));
        if (!(zoom && far < dlo - 2)) {
          // a slot touches the tile when its > tau ellipse  M[dr,dc] <= T^2  meets the tile's
          // rectangle (1.5 pixels of slack): the form's minimum over a box that does not hold
          // the centre lies on one of the four edges.  No square roots, no divisions.
          const float ra = (float)r0 - 1.5f, rb = (float)(rend - 1) + 1.5f;
          const float ca = (float)tc0 - 1.5f, cb_ = (float)(min(tc0 + kTileCols, N) - 1) + 1.5f;
          unsigned m = 0u;
#pragma unroll 4
          for (int pos = 0; pos < base; ++pos) {  // branch-free: the loads of a batch overlap
            const int2 bbr = *reinterpret_cast<const int2*>(&sm.coef[cb][pos].r0);  // r0 r1
            const int2 bbc = *reinterpret_cast<const int2*>(&sm.coef[cb][pos].c0);  // c0 c1
            const int4 bb = make_int4(bbr.x, bbr.y, bbc.x, bbc.y);
            const float4 gm = *reinterpret_cast<const float4*>(&geo[pos].Mrr);     // Mrr Mcc Mrc T2
            const float4 gx = *reinterpret_cast<const float4*>(&geo[pos].ec);      // ec x y det
            const float4 gi = *reinterpret_cast<const float4*>(&geo[pos].inv_cc);  // inv_cc inv_rr - -
            const bool inbox = !(bb.y < r0 || bb.x >= rend || bb.w < tc0 || bb.z >= tc0 + kTileCols);
            const float dra = ra - gx.z, drb = rb - gx.z, dca = ca - gx.y, dcb = cb_ - gx.y;
            // only the edges that face the centre can hold the minimum
            const bool use_c = dca > 0.f || dcb < 0.f, use_r = dra > 0.f || drb < 0.f;
            const float e = dca > 0.f ? dca : dcb, f = dra > 0.f ? dra : drb;
            const float tc_ = fminf(fmaxf(-gm.z * e * gi.y, dra), drb);  // column edge dc = e, dr free
            const float qc = (gm.x * tc_ + 2.f * gm.z * e) * tc_ + gm.y * e * e;
            const float tr_ = fminf(fmaxf(-gm.z * f * gi.x, dca), dcb);  // row edge dr = f, dc free
            const float qr = (gm.y * tr_ + 2.f * gm.z * f) * tr_ + gm.x * f * f;
            const float big = 3.0e38f;
            float qmin = fminf(use_c ? qc : big, use_r ? qr : big);
            if (!use_c && !use_r) qmin = 0.f;  // the centre lies inside the rectangle
            // (NaN from a degenerate slot counts as touching)
            if (inbox && !(qmin > gm.w)) m |= 1u << pos;
          }
          hits[h] = m;
          cost[h] = 1 + min(__popc(hits[h]), 30);
          tag[h] = (g << 8) | tc;
          among[h] = atomicAdd(&sm.hist[cost[h]], 1);  // position among the tiles of equal cost
        }
      }
    }
    __syncwarp();
    // rank = tiles of higher cost + position among equals
    int above[TPL];
#pragma unroll
    for (int h = 0; h < TPL; ++h) above[h] = 0;
    for (int c = 31; c >= 1; --c) {
      const int n = sm.hist[c];
#pragma unroll
      for (int h = 0; h < TPL; ++h)
        if (c > cost[h]) above[h] += n;
    }
    // Snake deal over the warps without a side job: first_worker .. kWarps - 1.
    const int seats = kWarps - first_worker;
#pragma unroll
    for (int h = 0; h < TPL; ++h) {
      if (cost[h] == 0) continue;
      const int i = above[h] + among[h];
      int ps = (i >= seats) + (i >= 2 * seats) + (i >= 3 * seats) + (i >= 4 * seats) + (i >= 5 * seats);
      if constexpr (TPL > 2) ps += (i >= 6 * seats) + (i >= 7 * seats);
      const int j = i - ps * seats;
      const int w = first_worker + ((ps & 1) ? j : seats - 1 - j), e = ps;
      sm.rec[cb][w][e] = make_uint2(hits[h], (unsigned)tag[h]);
    }
  }
  RSB_BSTAMP(4);
}

// ---- float64 peak (jet_max) for the labels, one warp -------------------------------
// The float32 image fixes WHERE the maximum is; its float64 value is recomputed from the
// parameters: the raw pixels under the non-zero taps (<= 3 x 3) x slots, operation order of
// gauss.py:77-88 except that the two divisions by sigma are multiplications by per-slot
// reciprocals (a last-bit difference in the exponent's argument), combined with float64 tap
// weights.  scratch: >= 16*slots + 6*slots + 8 doubles.
__device__ inline double peak64(double* scratch, const double* P, int slots, int N, bool zoom,
                                int zpad, int margin, unsigned long long image_best, int lane) {
  const unsigned idx = ~(unsigned)(image_best & 0xffffffffull);
  const int pr = (int)(idx / (unsigned)N), pc = (int)(idx % (unsigned)N);
  double* rot = scratch + 16 * slots;      // per slot: cr, sr, ra, rb, 1/sy, 1/sx
  int* coord = reinterpret_cast<int*>(rot + 6 * slots);  // [0..3] rows, [4..7] columns of the non-zero taps
  for (int k = lane; k < slots; k += 32) {
    const double* p = P + 6 * k;
    double cr, sr;
    sincos(p[5], &sr, &cr);
    rot[6 * k] = cr;
    rot[6 * k + 1] = sr;
    rot[6 * k + 2] = p[2] * cr - p[1] * sr;
    rot[6 * k + 3] = p[2] * sr + p[1] * cr;
    rot[6 * k + 4] = 1.0 / p[4];
    rot[6 * k + 5] = 1.0 / p[3];
  }
  // taps with a non-zero weight, compacted: rows in even lanes' registers, columns in odd ones
  // (both halves of the warp work at the same time)
  Taps64 t;
  if (zoom) {
    t = composite_taps((lane & 1) ? pc : pr, N, margin, zpad);
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      t.i[k] = (lane & 1) ? pc : pr;
      t.w[k] = (k == 0) ? 1.0 : 0.0;
    }
  }
  double w[4] = {0.0, 0.0, 0.0, 0.0};
  int n = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (t.w[k] != 0.0) {
      // (n <= k: a chain of selects instead of a dynamically indexed register array)
#pragma unroll
      for (int m = 0; m < 4; ++m)
        if (m == n) w[m] = t.w[k];
      if (lane < 2) coord[4 * lane + n] = t.i[k];
      ++n;
    }
  }
  const int nr = __shfl_sync(0xffffffffu, n, 0), nc = __shfl_sync(0xffffffffu, n, 1);
  __syncwarp();
  const int npx = nr * nc;
  const unsigned inv_slots = (65536u + (unsigned)slots - 1u) / (unsigned)slots;  // exact below 1024
  for (int e = lane; e < npx * slots; e += 32) {
    const int px = (int)(((unsigned)e * inv_slots) >> 16), k = e - px * slots;
    const int a = (px >= nc) + (px >= 2 * nc) + (px >= 3 * nc), b = px - a * nc;
    const double* p = P + 6 * k;
    double v = 0.0;
    if (p[0] != 0.0) {
      const int r = coord[a], c = coord[4 + b];
      const double* ro = rot + 6 * k;
      const double ua = r * ro[0] - c * ro[1], ub = r * ro[1] + c * ro[0];
      const double ea = (ro[2] - ua) * ro[4], eb = (ro[3] - ub) * ro[5];
      v = p[0] * exp(-(ea * ea + eb * eb) / 2.0);
    }
    scratch[e] = v;
  }
  __syncwarp();
  double s = 0.0;
  if (lane < npx)
    for (int k = 0; k < slots; ++k) s += scratch[lane * slots + k];  // slot order
  // rows' weights live in lane 0, columns' in lane 1
  double acc = 0.0;
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    double h = 0.0;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const double wc = __shfl_sync(0xffffffffu, w[b], 1);
      const double sv = __shfl_sync(0xffffffffu, s, (a * nc + b) & 31);
      if (a < nr && b < nc) h += wc * sv;
    }
    const double wr = __shfl_sync(0xffffffffu, w[a], 0);
    if (a < nr) acc += wr * h;
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

// kVar: every image has its own size (a.sizes[img] <= a.size) and output offset -- the sub-jets
// of a survey (reference survey/survey.py:54-60) in ONE launch.  With two CTAs per image the
// first band takes ceil(N/2) rows, so odd sizes work: the centre row (N-1)/2 is the last row of
// the first band and the halo row of the second, and no tap with a non-zero weight reaches
// further across (the composite operator preserves the centre).
// kBig (one CTA per image only): the whole image of up to 224 rows sits in this CTA's shared memory.  The V
// pass cannot keep that many rows in registers, so it runs twice -- once for the maximum, once to scale and
// store -- which costs less than the second CTA's barriers and exchanges for images of 129 .. 224 pixels.
template <int kCluster, int NS, bool kVar = false, bool kBig = false>
__global__ void __launch_bounds__(kThreads, 1) jet2_kernel(const Args a) {
  static_assert(!kBig || kCluster == 1, "kBig is a single-CTA variant");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Fixed2& sm = *reinterpret_cast<Fixed2*>(smem_raw);
  const int slots = a.slots;
  const int npad_max = (a.size + 3) & ~3;
  // composite taps, one record per output coordinate: {w0, w1, w2, first tap (absolute, as int bits)}
  float4* tab = reinterpret_cast<float4*>(smem_raw + sizeof(Fixed2));  // [npad_max]
  float* raw = reinterpret_cast<float*>(tab + npad_max);               // [H (+1 halo)][pitch]

  const uint32_t s_raw = keep_u32(smem_u32(raw)), s_tab = keep_u32(smem_u32(tab));
  const int tid = keep_s32((int)threadIdx.x), lane = tid & 31, warp = tid >> 5;
  const int rank = keep_s32((kCluster > 1) ? (int)(blockIdx.x % kCluster) : 0);
  const int cid = blockIdx.x / kCluster, nclusters = gridDim.x / kCluster;
  const bool f64 = (a.flags & RSB_JET_OUT_F64) != 0;
  const bool normalize = (a.flags & RSB_JET_NORMALIZE) != 0;
  const size_t esz = f64 ? 8 : 4;
  // ---- geometry of the current image (fixed for the whole launch unless kVar) ----
  int N, npad, pitch, opitch;  // size, size rounded up to 4, raw row pitch, output row pitch (elements)
  int H0, H, R0, base;         // rows of the first band / of this CTA's band, its first absolute row,
                               // absolute row stored at raw row 0 (second band: the halo row)
  int strips, tcols, stage_rows;
  bool vec;
  float* stage;
  uint32_t s_stage;
  // the stage sits behind the largest raw band of the launch, so that an image never renders
  // into rows the previous image's bulk store is still reading
  const int Hmax = (kCluster == 1) ? a.size : (a.size + 1) / 2;
  float* const stage_var = raw + (size_t)(Hmax + (kCluster > 1 ? 1 : 0)) * npad_max;
  auto geometry = [&](int n_) {
    N = n_;
    npad = (N + 3) & ~3;
    pitch = kVar ? npad : a.pitch;
    opitch = kVar ? npad : N;
    H0 = (kCluster == 1) ? N : (N + 1) / 2;
    H = (kCluster == 1 || rank == 0) ? H0 : N - H0;
    R0 = keep_s32((rank == 0) ? 0 : H0);
    base = keep_s32((rank == 0) ? 0 : H0 - 1);
    strips = (N + kStrip - 1) / kStrip;           // 128-column strips (H/V pass)
    tcols = (N + kTileCols - 1) / kTileCols;      // 64-column tiles (phase 1)
    vec = (opitch & 3) == 0 && (reinterpret_cast<uintptr_t>(a.out) & 15) == 0;
    // float32 output with contiguous rows: the first stage_rows rows of the finished band are
    // staged in shared memory and leave through one bulk async copy that drains while the next
    // image is being rendered (an SM pushes ~30 B/clk towards L2 whichever engine stores, so
    // the gain is the overlap, not the engine); the other rows are stored directly.
    if constexpr (kVar) {
      const int room = (a.raw_floats - (int)(stage_var - raw)) / opitch;
      stage_rows = (!f64 && vec) ? max(0, min(H / 2, room)) : 0;
      stage = stage_var;
    } else {
      stage_rows = (!f64 && vec && pitch == N) ? a.stage_rows : 0;
      stage = raw + (size_t)(H + (kCluster > 1 ? 1 : 0)) * pitch;
    }
    s_stage = keep_u32(smem_u32(stage));
  };
  geometry((kVar && cid < a.n) ? a.sizes[cid] : a.size);
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
  const char* bulk_dst = nullptr;  // (thread 0) where the staged rows of the previous image go
  uint32_t bulk_bytes = 0;         // ... and how many bytes of the stage they are

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
    sm.next_row = 0;
  }
  __syncthreads();
  // ---- which images this cluster renders --------------------------------------------------------------
  // A zoom image costs ~1.8 x a plain one, so striding through the batch gives the clusters unequal sums (the
  // slowest of 74 ends ~3 % behind the mean at 16 384 images).  Instead the zoom images of the launch are dealt
  // round-robin to the clusters, and the plain ones round-robin from the other end: every cluster gets the same
  // number of each (+- 1).  Both CTAs of a cluster build the same list.
  // (the launcher keeps a.n <= kListCap x clusters; everything below the list lives in shared memory, not in
  // registers: the render loops have none to spare)
  {
    const int per = ((a.n + kWarps * 32 - 1) / (kWarps * 32)) * 32;  // images per warp (a multiple of 32)
    const int lo = min(a.n, warp * per), hi = min(a.n, lo + per);
    int zc = 0, pc = 0;
    for (int b0 = lo; b0 < hi; b0 += 32) {
      const int i = b0 + lane;
      const bool v = i < hi, z = v && a.zoom_pad[i] >= 0;
      zc += __popc(__ballot_sync(0xffffffffu, z));
      pc += __popc(__ballot_sync(0xffffffffu, v && !z));
    }
    if (lane == 0) {
      sm.wz[warp] = zc;
      sm.wp[warp] = pc;
    }
    __syncthreads();
    int z0 = 0, p0 = 0, Z = 0, P = 0;
    for (int w = 0; w < kWarps; ++w) {
      if (w < warp) { z0 += sm.wz[w]; p0 += sm.wp[w]; }
      Z += sm.wz[w];
      P += sm.wp[w];
    }
    const int back = nclusters - 1 - cid;  // the plain images are dealt from the other end
    const int nz = Z > cid ? (Z - cid + nclusters - 1) / nclusters : 0;
    if (tid == 0) sm.nimg = nz + (P > back ? (P - back + nclusters - 1) / nclusters : 0);
    const unsigned lt = (1u << lane) - 1u;
    for (int b0 = lo; b0 < hi; b0 += 32) {
      const int i = b0 + lane;
      const bool v = i < hi, z = v && a.zoom_pad[i] >= 0;
      const unsigned mz = __ballot_sync(0xffffffffu, z), mp = __ballot_sync(0xffffffffu, v && !z);
      if (z) {
        const int r = z0 + __popc(mz & lt);
        if (r % nclusters == cid) sm.img_list[r / nclusters] = i;
      } else if (v) {
        const int r = p0 + __popc(mp & lt);
        if (r % nclusters == back) sm.img_list[nz + r / nclusters] = i;   // (zoom ones first, then the plain ones)
      }
      z0 += __popc(mz);
      p0 += __popc(mp);
    }
    __syncthreads();
  }
  const uint32_t s_list = smem_u32(&sm.img_list[0]), s_nimg = smem_u32(&sm.nimg);
  auto image_at = [&](int k) -> int { return lds_s32(s_list + (uint32_t)k * 4u); };  // k < images()
  auto images = [&]() -> int { return lds_s32(s_nimg); };
  auto prefetch = [&](int image, int pb) {  // one thread: parameter block -> params[pb]
    fence_proxy_async();
    mbar_expect_tx(&sm.bar[pb], pbytes);
    bulk_g2s(sm.params[pb], a.params + (size_t)image * slots * 6, pbytes, &sm.bar[pb]);
  };
  const int img0 = images() > 0 ? image_at(0) : 0;
  if constexpr (kVar) {
    if (images() > 0) geometry(a.sizes[img0]);  // (the first image is no longer image number cid)
  }
  if (tid == 0 && images() > 0) {
    prefetch(img0, 0);
    if (images() > 1) prefetch(image_at(1), 1);
  }
  // warp 1 only has its side job (the labels) in the cluster's first CTA; elsewhere it renders
  const int first_worker = (rank == 0) ? 2 : 1;
  if (warp == 0 && images() > 0) {  // prologue: coefficients of the first image
    mbar_wait(&sm.bar[0], 0);
    if constexpr (kBig) build_coefs2_big<4>(sm, 0, sm.params[0], slots, N, R0, H, lane, a.zoom_pad[img0] >= 0, first_worker);
    else build_coefs2(sm, 0, sm.params[0], slots, N, R0, H, lane, a.zoom_pad[img0] >= 0, first_worker);
  }
  // the peer CTA is resident and its mbarriers are initialised before any st.async
  if constexpr (kCluster > 1) cg::this_cluster().sync(); else __syncthreads();
  int zpad_next = (images() > 0) ? a.zoom_pad[img0] : -1;

  // Two warps start each image with a side job before they join the tile queue: warp 0 builds
  // the NEXT image's coefficients and culling spans, warp 1 turns the PREVIOUS image's arg-max
  // into the float64 peak label.  (Dedicated helper warps were tried: no gain, more spills.)
  constexpr int kCoefWarp = 0, kLabelWarp = 1;
  auto worker_sync = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(kWorkers) : "memory"); };
  auto label_job = [&](int par) {
    const Fixed2::LabelReq q = sm.label[par];
    double peak = 1.0;
    if (normalize)
      peak = peak64(sm.pix64, sm.params[q.pb], slots, q.n, q.zpad >= 0, q.zpad, q.margin, q.best, lane);
    if (lane == 0) {
      a.margin[q.img] = q.margin;
      a.peak[q.img] = peak;
    }
  };

  int it = 0;
  for (; it < images(); ++it) {
    const int img = image_at(it);
    const int img_next = (it + 1 < images()) ? image_at(it + 1) : -1;  // (-1: this was the last one)
    const int buf = it & 1;   // coefficient / exchange-slot parity
    const int pb = it % 3;    // parameter buffer
    mbar_wait(&sm.bar[pb], (it / 3) & 1);
    __syncthreads();  // (A) coef/span/nlive of this image visible; raw free for reuse; stage complete
    if (tid == 0 && bulk_dst) {
      bulk_s2g(const_cast<char*>(bulk_dst), stage, bulk_bytes);
      bulk_dst = nullptr;
    }
    if constexpr (kVar) {
      if (it > 0) geometry(a.sizes[img]);
    }
    if (tid == 0) {  // running reductions of this image (all reads of the previous values are behind A)
      sm.lbest = 0ull;
      sm.lmargin = N;
      sm.ncand = 0;
      sm.next_row = 0;
      sm.foot_lo = N;
      sm.foot_hi = -1;
      sm.scale[0] = (double)(N + 2 * max(zpad_next, 0)) / (double)N;  // (zpad_next: this image's pad)
    }
    __syncwarp();
    {
      if (warp == kCoefWarp) {
        if (img_next >= 0) {
          const int nb = (it + 1) % 3;
          mbar_wait(&sm.bar[nb], ((it + 1) / 3) & 1);
          int nN = N, nR0 = R0, nH = H;  // the next image's geometry
          if constexpr (kVar) {
            nN = a.sizes[img_next];
            const int g0 = (kCluster == 1) ? nN : (nN + 1) / 2;
            nH = (kCluster == 1 || rank == 0) ? g0 : nN - g0;
            nR0 = (rank == 0) ? 0 : g0;
          }
#ifdef RSB_TIMELINE
          build_coefs2(sm, buf ^ 1, sm.params[nb], slots, nN, nR0, nH, lane, a.zoom_pad[img_next] >= 0,
                       first_worker, g_timeline ? &g_timeline[((size_t)img * 2 + rank) * 96 + 24] : nullptr);
#else
          if constexpr (kBig)
            build_coefs2_big<4>(sm, buf ^ 1, sm.params[nb], slots, nN, nR0, nH, lane,
                                a.zoom_pad[img_next] >= 0, first_worker);
          else
            build_coefs2(sm, buf ^ 1, sm.params[nb], slots, nN, nR0, nH, lane,
                         a.zoom_pad[img_next] >= 0, first_worker);
#endif
        }
        RSB_STAMP_T(12, kCoefWarp * 32);
      } else if (warp == kLabelWarp) {
        // previous image's labels, then its parameter buffer is free for image it+2
        if (rank == 0 && it > 0) label_job(buf ^ 1);
        if (lane == 0 && it + 2 < images()) prefetch(image_at(it + 2), (it + 2) % 3);
        RSB_STAMP_T(13, kLabelWarp * 32);
      }
    }
    RSB_STAMP(0);
    const double* P = sm.params[pb];
    const int zpad = zpad_next;
    if (img_next >= 0) zpad_next = a.zoom_pad[img_next];  // latency hidden
    const bool zoom = zpad >= 0;
    const float thr = (float)kEdgeThr, delta = sm.delta[buf];
    const uint32_t s_coef = keep_u32(smem_u32(sm.coef[buf]));
    const uint32_t pitch4 = keep_u32((uint32_t)pitch * 4u);

    // ---------------- phase 1: splat this CTA's band, 8 x 64 tiles ------------------
    // Each warp walks its own work list (built one image ahead by build_coefs2).  Everything
    // that comes out of shared memory is requested one step before it is needed -- the next
    // tile's record during this tile, the next slot's coefficients during this slot -- because
    // a shared-memory round trip queues behind the other warps' MUFU work.
    int my_margin = N;
    unsigned long long my_best = 0ull;
    float best_v = -1.f;  // this lane's running maximum and its pixel index
    int best_i = 0;
#ifdef RSB_TIMELINE
    long long dbg_t0 = clock64(), dbg_hit = 0, dbg_epi = 0, dbg_ntile = 0, dbg_nhit = 0;
    const long long dbg_start = dbg_t0;
#define DBG_LAP(acc) do { const long long _n = clock64(); acc += _n - dbg_t0; dbg_t0 = _n; } while (0)
#else
#define DBG_LAP(acc) do {} while (0)
#endif
    const uint32_t s_rec = keep_u32(smem_u32(&sm.rec[buf][warp][0]));
    const float tl = thr - delta, th = thr + delta;
    uint2 rc = lds_v2u(s_rec);
    // coefficients of the slot about to be rendered
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f), ct = q;
    float2 kk = make_float2(0.f, 0.f);
    auto load_first = [&](const uint2& r) {
      if (r.y != kRecEnd && r.x) {
        const uint32_t an = s_coef + (uint32_t)lowest_bit(r.x) * 64u;
        q = lds_v4(an);         // A B C la
        ct = lds_v4(an + 16u);  // irf fr icf fc
        kk = lds_v2(an + 32u);  // k2 k3
      }
    };
    load_first(rc);
    for (int e = 0; rc.y != kRecEnd; ++e) {
      uint2 nx = make_uint2(0u, kRecEnd);
      if (e + 1 < kMaxRec) nx = lds_v2u(s_rec + (uint32_t)(e + 1) * 8u);
      const int g = (int)(rc.y >> 8), tc = (int)(rc.y & 0xffu);
      const int r0 = R0 + g * kTileRows, tc0 = tc * kTileCols, c0 = tc0 + 2 * lane;
      const uint32_t s_dst = s_raw + (uint32_t)((r0 - base) * pitch + c0) * 4u;
      const int rend = min(R0 + H, r0 + kTileRows);  // rows of this tile that exist
      const float r0f = (float)r0, c0f = (float)c0;
#ifdef RSB_TIMELINE
      dbg_ntile++;
      dbg_nhit += __popc(rc.x);
#endif
      if (rc.x == 0u) {
        // no slot touches this tile: zeros (the resampling taps may still read them)
        load_first(nx);
        if (c0 < N) {
#pragma unroll
          for (int i = 0; i < kTileRows; ++i)
            if (r0 + i < rend) sts_b64(s_dst + (uint32_t)i * pitch4, 0ull);
        }
        rc = nx;
        continue;
      }
      f2 acc2[kTileRows];  // two adjacent pixels per row, packed (FFMA2)
#pragma unroll
      for (int i = 0; i < kTileRows; ++i) acc2[i] = 0ull;
      // one slot over this lane's 16 pixels of the tile
      auto slot_math = [&](const float4& cq, const float4& cc, const float2& ck) {
        const float d0 = (c0f - cc.z) - cc.w;
        const f2 d = f2_pack(d0, d0 + 1.f);
        const f2 Q = f2_mul(f2_pack(cq.y, cq.y), d);
        const float dr0 = (r0f - cc.x) - cc.y;
        if (ck.x > 0.f) {
          // rows in two blocks of four, two exponentials per column and block (see Coef2)
          const f2 Pq = f2_fma(f2_mul(f2_pack(cq.x, cq.x), d), d, f2_pack(cq.w, cq.w));  // A dc^2 + la
          const f2 k2 = f2_pack(ck.x, ck.x), k3 = f2_pack(ck.y, ck.y);
#pragma unroll
          for (int b = 0; b < kTileRows; b += 4) {
            const float dr = dr0 + (float)b;
            const float Cq = cq.z * dr * dr;
            const float tw = fmaf(2.f * cq.z, dr, cq.z);  // 2 C dr + C
            float ta, tb, wa, wb;
            f2_unpack(f2_fma(f2_pack(dr, dr), Q, f2_add(Pq, f2_pack(Cq, Cq))), ta, tb);
            f2_unpack(f2_add(Q, f2_pack(tw, tw)), wa, wb);
            // (the clamp never binds where the block holds a pixel above the cut-off, sigma >=
            // 0.6: there |w| < 31; elsewhere it keeps U^3 finite)
            const f2 E0 = f2_pack(P1_EX2(ta), P1_EX2(tb));
            const f2 U = f2_pack(P1_EX2(fminf(wa, 40.f)), P1_EX2(fminf(wb, 40.f)));
            const f2 U2 = f2_mul(U, U);
            acc2[b] = f2_add(acc2[b], E0);
            acc2[b + 1] = f2_fma(E0, U, acc2[b + 1]);
            acc2[b + 2] = f2_fma(E0, f2_mul(U2, k2), acc2[b + 2]);
            acc2[b + 3] = f2_fma(E0, f2_mul(f2_mul(U2, U), k3), acc2[b + 3]);
          }
        } else {
          // one exponential per pixel; cq.w holds the amplitude itself
          const f2 Pq = f2_mul(f2_mul(f2_pack(cq.x, cq.x), d), d);
          const f2 amp2 = f2_pack(cq.w, cq.w);
#pragma unroll
          for (int i = 0; i < kTileRows; i += 2) {
            // two rows at a time: dr and C dr^2 as a packed pair, then broadcast per row
            const f2 drp = f2_pack(dr0 + (float)i, dr0 + (float)(i + 1));
            const f2 Cqp = f2_mul(f2_mul(f2_pack(cq.z, cq.z), drp), drp);
            float dra, drb, Cqa, Cqb;
            f2_unpack(drp, dra, drb);
            f2_unpack(Cqp, Cqa, Cqb);
            float ta, tb;
            f2_unpack(f2_fma(f2_pack(dra, dra), Q, f2_add(Pq, f2_pack(Cqa, Cqa))), ta, tb);
            acc2[i] = f2_fma(amp2, f2_pack(P1_EX2(ta), P1_EX2(tb)), acc2[i]);
            f2_unpack(f2_fma(f2_pack(drb, drb), Q, f2_add(Pq, f2_pack(Cqb, Cqb))), ta, tb);
            acc2[i + 1] = f2_fma(amp2, f2_pack(P1_EX2(ta), P1_EX2(tb)), acc2[i + 1]);
          }
        }
      };
      // The next slot's coefficients are requested before this slot's arithmetic (slot 0 when
      // there is none: always valid memory, never used).  The warp barrier pins the loads above
      // the arithmetic: ptxas otherwise sinks them to the end of the iteration, where their
      // latency is exposed.  Unrolled by two so that the two coefficient sets swap roles
      // without register moves.
      unsigned rest = rc.x;  // slots still to render; the lowest one's coefficients are in q, ct, kk
      float4 qb, cb;
      float2 kb;
#ifndef RSB_NO_PAIR
      // Two slots per step: their arithmetic is one straight-line block (both of the four-row-block kind),
      // so the scheduler has two independent dependency chains per warp to draw from -- with 4 warps per
      // scheduler the splat is bound by each warp's own chains, not by issue slots.
      auto blocks_math = [&](const float4& cq, const float4& cc, const float2& ck) {
        const float d0 = (c0f - cc.z) - cc.w;
        const f2 d = f2_pack(d0, d0 + 1.f);
        const f2 Q = f2_mul(f2_pack(cq.y, cq.y), d);
        const float dr0 = (r0f - cc.x) - cc.y;
        const f2 Pq = f2_fma(f2_mul(f2_pack(cq.x, cq.x), d), d, f2_pack(cq.w, cq.w));
        const f2 k2 = f2_pack(ck.x, ck.x), k3 = f2_pack(ck.y, ck.y);
#pragma unroll
        for (int b = 0; b < kTileRows; b += 4) {
          const float dr = dr0 + (float)b;
          const float Cq = cq.z * dr * dr;
          const float tw = fmaf(2.f * cq.z, dr, cq.z);
          float ta, tb, wa, wb;
          f2_unpack(f2_fma(f2_pack(dr, dr), Q, f2_add(Pq, f2_pack(Cq, Cq))), ta, tb);
          f2_unpack(f2_add(Q, f2_pack(tw, tw)), wa, wb);
          const f2 E0 = f2_pack(P1_EX2(ta), P1_EX2(tb));
          const f2 U = f2_pack(P1_EX2(fminf(wa, 40.f)), P1_EX2(fminf(wb, 40.f)));
          const f2 U2 = f2_mul(U, U);
          acc2[b] = f2_add(acc2[b], E0);
          acc2[b + 1] = f2_fma(E0, U, acc2[b + 1]);
          acc2[b + 2] = f2_fma(E0, f2_mul(U2, k2), acc2[b + 2]);
          acc2[b + 3] = f2_fma(E0, f2_mul(f2_mul(U2, U), k3), acc2[b + 3]);
        }
      };
      while (rest) {
        rest &= rest - 1;             // (q, ct, kk) hold the slot just removed
        if (!rest) {
          __syncwarp();
          slot_math(q, ct, kk);
          break;
        }
        {
          const uint32_t an = s_coef + (uint32_t)lowest_bit(rest) * 64u;
          qb = lds_v4(an);
          cb = lds_v4(an + 16u);
          kb = lds_v2(an + 32u);
        }
        rest &= rest - 1;
        const float4 qa = q, ca_ = ct;
        const float2 ka = kk;
        {
          const uint32_t an = s_coef + (uint32_t)max(lowest_bit(rest), 0) * 64u;   // next step's first slot
          q = lds_v4(an);
          ct = lds_v4(an + 16u);
          kk = lds_v2(an + 32u);
        }
        __syncwarp();
        if (ka.x > 0.f && kb.x > 0.f) {
          blocks_math(qa, ca_, ka);
          blocks_math(qb, cb, kb);
        } else {
          slot_math(qa, ca_, ka);
          slot_math(qb, cb, kb);
        }
      }
#else
      while (rest) {
        rest &= rest - 1;
        {
          const uint32_t an = s_coef + (uint32_t)max(lowest_bit(rest), 0) * 64u;
          qb = lds_v4(an);
          cb = lds_v4(an + 16u);
          kb = lds_v2(an + 32u);
        }
        __syncwarp();
        slot_math(q, ct, kk);
        if (!rest) break;
        rest &= rest - 1;
        {
          const uint32_t an = s_coef + (uint32_t)max(lowest_bit(rest), 0) * 64u;
          q = lds_v4(an);
          ct = lds_v4(an + 16u);
          kk = lds_v2(an + 32u);
        }
        __syncwarp();
        slot_math(qb, cb, kb);
      }
#endif
      load_first(nx);  // in flight during the epilogue
      if (c0 + 1 >= N) {  // odd size: the pair's second column lies outside the image
#pragma unroll
        for (int i = 0; i < kTileRows; ++i) {
          float lo_, hi_;
          f2_unpack(acc2[i], lo_, hi_);
          acc2[i] = f2_pack(lo_, 0.f);
        }
      }
      float acc[kTileRows][2];
#pragma unroll
      for (int i = 0; i < kTileRows; ++i) f2_unpack(acc2[i], acc[i][0], acc[i][1]);
      DBG_LAP(dbg_hit);
      if (c0 < N) {
        float tile_top = 0.f;
        const bool full = r0 + kTileRows <= rend;  // always, when the band is a multiple of 8 rows
        if (full) {
#pragma unroll
          for (int i = 0; i < kTileRows; ++i) {
            sts_b64(s_dst + (uint32_t)i * pitch4, acc2[i]);
            tile_top = fmaxf(tile_top, fmaxf(acc[i][0], acc[i][1]));
          }
        } else {
#pragma unroll
          for (int i = 0; i < kTileRows; ++i) {
            if (r0 + i < rend) {
              sts_b64(s_dst + (uint32_t)i * pitch4, acc2[i]);
              tile_top = fmaxf(tile_top, fmaxf(acc[i][0], acc[i][1]));
            }
          }
        }
        if (zoom) {
          // border distance of this lane's nearest-to-border pixel in the tile
          const int rl = rend - 1;
          const int dmin = min(min(r0, N - 1 - rl), min(c0, N - 1 - min(c0 + 1, N - 1)));
          if (tile_top > tl && dmin < my_margin) {
            // The border distance of pixel (r, c) is min(db(r), dc(c)), so its minimum over the
            // pixels above the threshold splits into "first row with a pixel above" and "first
            // column with a pixel above" -- when the tile lies on one side of both centre
            // lines, rows and columns are monotone in border distance.
            const bool top = 2 * (r0 + kTileRows - 1) <= N - 1, bottom = 2 * r0 >= N - 1;
            const bool left = 2 * (c0 + 1) <= N - 1, right = 2 * c0 >= N - 1;
            bool done = false;
            if (full && c0 + 1 < N && (top || bottom) && (left || right)) {
              float rowmax[kTileRows], colmax[2] = {0.f, 0.f};
#pragma unroll
              for (int i = 0; i < kTileRows; ++i) {
                rowmax[i] = fmaxf(acc[i][0], acc[i][1]);
                colmax[0] = fmaxf(colmax[0], acc[i][0]);
                colmax[1] = fmaxf(colmax[1], acc[i][1]);
              }
              const int dc_[2] = {left ? c0 : N - 1 - c0, left ? c0 + 1 : N - 2 - c0};
              auto nearest = [&](float level, bool strict) -> int {
                int d = 0x7fffffff;
#pragma unroll
                for (int i = 0; i < kTileRows; ++i) {
                  const bool on = strict ? rowmax[i] > level : rowmax[i] >= level;
                  const int db = top ? r0 + i : N - 1 - (r0 + i);
                  if (on) d = min(d, db);
                }
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                  const bool on = strict ? colmax[j] > level : colmax[j] >= level;
                  if (on) d = min(d, dc_[j]);
                }
                return d;
              };
              const int d_lo = nearest(tl, true);
              if (d_lo >= my_margin) {
                done = true;  // nothing here can lower the margin
              } else if (nearest(th, false) == d_lo) {
                my_margin = d_lo;  // reached by a pixel definitely above the threshold
                done = true;
              }
            }
            if (!done) {
#pragma unroll
              for (int i = 0; i < kTileRows; ++i) {
                const int r = r0 + i;
                if (r >= rend) break;
                const int db = min(r, N - 1 - r);
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                  const int c = c0 + j;
                  const int dd = min(db, min(c, N - 1 - c));
                  if (c < N && dd < my_margin && acc[i][j] > tl) {
                    if (acc[i][j] >= th) {
                      my_margin = dd;  // definitely above
                    } else {
                      // within float32 noise of the threshold: float64 decides, later and only
                      // if it can still lower the margin
                      const int slot = atomicAdd(&sm.ncand, 1);
                      if (slot < kMaxCand) sm.cand[slot] = (r << 16) | c;
                      else if (eval64(P, slots, r, c) > kEdgeThr) my_margin = dd;
                    }
                  }
                }
              }
            }
          }
        } else if (normalize && tile_top >= best_v) {
          // first pixel (smallest index) that holds the tile's maximum; tiles come in no
          // particular order, so equal values are decided by the index
          int pi = 0x7fffffff;
#pragma unroll
          for (int i = kTileRows - 1; i >= 0; --i) {
            if (r0 + i < rend) {
              if (acc[i][1] == tile_top) pi = (r0 + i) * N + c0 + 1;
              if (acc[i][0] == tile_top) pi = (r0 + i) * N + c0;
            }
          }
          if (tile_top > best_v || pi < best_i) {
            best_v = tile_top;
            best_i = pi;
          }
        }
      }
      DBG_LAP(dbg_epi);
      rc = nx;
    }
#ifdef RSB_TIMELINE
    if (g_timeline && tid == 64) {
      long long* tlp = &g_timeline[((size_t)img * 2 + rank) * 96 + 16];
      tlp[1] = dbg_hit; tlp[2] = dbg_epi; tlp[3] = dbg_ntile; tlp[4] = dbg_nhit;
    }
    if (g_timeline && lane == 0) {
      long long* tlp = &g_timeline[((size_t)img * 2 + rank) * 96 + 32 + warp * 4];
      tlp[0] = clock64(); tlp[1] = dbg_ntile; tlp[2] = dbg_nhit; tlp[3] = dbg_start;
    }
#endif

    RSB_STAMP(1);  // this warp is out of tiles
    int margin = 0;
    char* out = reinterpret_cast<char*>(a.out) +
                (kVar ? (size_t)a.out_offs[img] : (size_t)img * N * N) * esz;
    unsigned long long image_best = 0ull;  // arg-max key of the final image (normalize only)

    if (!zoom) {
      // ------------- plain branch ---------------------------------------------------
      if (normalize) {
        my_best = warp_max_u64(best_v >= 0.f ? best_key(best_v, best_i) : 0ull);
        if (lane == 0) atomicMax(&sm.lbest, my_best);
      }
      if (tid == 0) bulk_wait_read();  // the previous image's staged rows have left the stage
      worker_sync();
      image_best = sm.lbest;
      if (normalize) {
        const unsigned long long theirs = exchange(2, image_best);
        image_best = theirs > image_best ? theirs : image_best;
      }
      RSB_STAMP(2);
      const float inv = normalize ? 1.f / __uint_as_float((unsigned)(image_best >> 32)) : 1.f;
      const int q4 = npad >> 2;
      if (!f64 && vec && pitch == opitch) {
        // float32, 16-byte aligned contiguous rows: the band is one run of float4
        const uint32_t s_src = s_raw + (uint32_t)((R0 - base) * pitch) * 4u;
        float4* outv = reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + (size_t)R0 * opitch);
        const int nstage = stage_rows * q4;
        for (int e = tid; e < H * q4; e += kWorkers) {
          float4 v = lds_v4(s_src + (uint32_t)e * 16u);
          v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
          if (e < nstage) sts_v4(s_stage + (uint32_t)e * 16u, v);
          else outv[e] = v;
        }
      } else {
        for (int e = tid; e < H * q4; e += kWorkers) {
          const int rl = e / q4, c0 = (e - rl * q4) << 2;
          const int r = R0 + rl;
          float4 v = lds_v4(s_raw + (uint32_t)((r - base) * pitch + c0) * 4u);
          v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
          store_out4(out + (size_t)r * opitch * esz, f64, c0, N, vec, v);
        }
      }
    } else {
      // ------------- zoom branch ------------------------------------------------------
      my_margin = warp_min(my_margin);
      if (lane == 0 && my_margin < N) atomicMin(&sm.lmargin, my_margin);
      worker_sync();
      {
        const int ncand = min(sm.ncand, kMaxCand);
#ifdef RSB_TIMELINE
        if (g_timeline && tid == 64) g_timeline[((size_t)img * 2 + rank) * 96 + 29] = sm.ncand;
#endif
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
                for (int l = 0; l < min(32, slots - k0); ++l) s += __shfl_sync(0xffffffffu, t, l);  // slot order
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
      // the two scale factors of the composite operator: the pad's was computed by thread 0 at
      // the start of the image (foot_* were reset there too), the crop's needs the margin
      const double sP = sm.scale[0], sC = (double)(N - 2 * margin) / (double)N;
      RSB_STAMP(30);

      // composite tap table (float64 math, float32 weights) + output footprint
      {
        int lo = N, hi = -1;
        for (int d = tid; d < npad; d += kWorkers) {
          Tap3 m;
          m.w[0] = m.w[1] = m.w[2] = 0.f;
          m.i = 0;
          if (d < N) {
            m = merge_taps(composite_taps(d, N, margin, zpad, sP, sC), N);
            if (m.w[0] != 0.f || m.w[1] != 0.f || m.w[2] != 0.f) {
              lo = min(lo, d);
              hi = max(hi, d);
            }
          }
          tab[d] = make_float4(m.w[0], m.w[1], m.w[2], __int_as_float(m.i));
        }
        RSB_STAMP(31);
        lo = __reduce_min_sync(0xffffffffu, lo);
        hi = __reduce_max_sync(0xffffffffu, hi);
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
            const float4 tp = lds_v4(s_tab + (uint32_t)c * 16u);
            const bool on = act[s] && cj < N;
            ct[s][j].w[0] = on ? tp.x : 0.f;
            ct[s][j].w[1] = on ? tp.y : 0.f;
            ct[s][j].w[2] = on ? tp.z : 0.f;
            ct[s][j].i = act[s] ? __float_as_int(tp.w) : max((N >> 1) - 1, 0);
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
      if (g_timeline && tid == 64) g_timeline[((size_t)img * 2 + rank) * 96 + 11] = dbg_rows_out;
#endif
      RSB_STAMP(4);
      worker_sync();  // (F) T complete in this band
      RSB_STAMP(5);
      if constexpr (kCluster > 1) {
        // halo: my boundary row of T goes into the peer's halo slot
        const int src = (rank == 0) ? (H - 1) : 1;  // local index of absolute row H0-1 / H0
        const int dst = (rank == 0) ? 0 : H0;       // peer's halo slot
        for (int e = tid; e < (pitch >> 2); e += kWorkers)
          st_async_v4(peer_raw + (uint32_t)(dst * pitch + 4 * e) * 4u,
                      reinterpret_cast<const float4*>(raw + (size_t)src * pitch)[e], peer_xbar + 8u);
        if (tid == 0) mbar_expect_tx(&sm.xbar[1], (uint32_t)pitch * 4u);
        mbar_wait(&sm.xbar[1], (uint32_t)(xuse[1] & 1));
        xuse[1]++;
      }

      RSB_STAMP(6);
      const int vr0 = max(R0, flo), vr1 = min(R0 + H - 1, fhi);  // inclusive footprint rows
      // local rows that hold T: the crop [margin, N - margin) within this band and its halo row
      const int tlo = max(margin - base, 0);
      const int thi = max(min(N - margin - 1 - base, (kCluster > 1) ? H : H - 1), tlo);
      if constexpr (kBig) {
        // V pass twice: the maximum first, then the same values again, scaled and stored
        auto vrow = [&](int r, int s) -> float4 {   // four output pixels of row r, strip s (zeros off the footprint)
          const int sc0 = s * kStrip, c0 = sc0 + 4 * lane;
          float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
          if (r < vr0 || r > vr1 || !(s < strips && sc0 <= fhi && sc0 + kStrip - 1 >= flo) || c0 >= N) return v;
          const float4 tp = lds_v4(s_tab + (uint32_t)r * 16u);
          const int l0 = __float_as_int(tp.w);
          // (zero-weight taps are clamped into the crop: only its rows hold T)
          const uint32_t p0 = s_raw + (uint32_t)(min(max(l0, tlo), thi) * pitch + c0) * 4u;
          const uint32_t p1 = s_raw + (uint32_t)(min(max(l0 + 1, tlo), thi) * pitch + c0) * 4u;
          const uint32_t p2 = s_raw + (uint32_t)(min(max(l0 + 2, tlo), thi) * pitch + c0) * 4u;
          const float4 t0 = lds_v4(p0), t1 = lds_v4(p1), t2 = lds_v4(p2);
          const f2 W0 = f2_pack(tp.x, tp.x), W1 = f2_pack(tp.y, tp.y), W2 = f2_pack(tp.z, tp.z);
          f2_unpack(f2_fma(W2, f2_pack(t2.x, t2.y), f2_fma(W1, f2_pack(t1.x, t1.y), f2_mul(W0, f2_pack(t0.x, t0.y)))), v.x, v.y);
          f2_unpack(f2_fma(W2, f2_pack(t2.z, t2.w), f2_fma(W1, f2_pack(t1.z, t1.w), f2_mul(W0, f2_pack(t0.z, t0.w)))), v.z, v.w);
          return v;
        };
        float inv = 1.f;
        if (normalize) {
          best_v = -1.f;
          best_i = 0;
          for (int r = vr0 + warp; r <= vr1; r += kWarps) {
#pragma unroll
            for (int s = 0; s < NS; ++s) {
              const int c0 = s * kStrip + 4 * lane;
              if (s >= strips || c0 >= N) continue;
              float4 m = vrow(r, s);
              if (c0 + 3 >= N) {  // ragged last group: columns >= N do not take part
                if (c0 + 1 >= N) m.y = -1.f;
                if (c0 + 2 >= N) m.z = -1.f;
                m.w = -1.f;
              }
              const float top = fmaxf(fmaxf(m.x, m.y), fmaxf(m.z, m.w));
              const int pi = r * N + c0 + (m.x == top ? 0 : (m.y == top ? 1 : (m.z == top ? 2 : 3)));
              if (top > best_v || (top == best_v && pi < best_i)) {
                best_v = top;
                best_i = pi;
              }
            }
          }
          my_best = warp_max_u64(best_v >= 0.f ? best_key(best_v, best_i) : 0ull);
          if (lane == 0) atomicMax(&sm.lbest, my_best);
          worker_sync();
          image_best = sm.lbest;
          inv = 1.f / __uint_as_float((unsigned)(image_best >> 32));
        }
        for (int r = R0 + warp; r < R0 + H; r += kWarps) {
#pragma unroll
          for (int s = 0; s < NS; ++s) {
            const int c0 = s * kStrip + 4 * lane;
            if (s >= strips || c0 >= N) continue;
            float4 v = vrow(r, s);
            v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
            store_out4(out + (size_t)r * opitch * esz, f64, c0, N, vec, v);
          }
        }
      } else {
      // V pass: results stay in registers (<= 8 rows x NS strips of 4 columns per lane)
      float4 res[kRowsPerWarp][NS];
      // running maximum of this lane: value and pixel index (its pixels come in increasing
      // index order, so '>' keeps the first of equal values)
      best_v = -1.f;
      best_i = 0;
#pragma unroll
      for (int i = 0; i < kRowsPerWarp; ++i) {
        const int r = R0 + warp + i * kWarps;
#pragma unroll
        for (int s = 0; s < NS; ++s) res[i][s] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < vr0 || r > vr1) continue;  // warp-uniform
        const float4 tp = lds_v4(s_tab + (uint32_t)r * 16u);
        const float w0 = tp.x, w1 = tp.y, w2 = tp.z;
        // only zero-weight taps can fall outside the crop or this band (+ halo): clamp them onto
        // rows that hold T (0 x whatever an unrendered tile left behind must not become NaN)
        const int l0 = __float_as_int(tp.w) - base;
        const uint32_t p0 = s_raw + (uint32_t)(min(max(l0, tlo), thi) * pitch) * 4u;
        const uint32_t p1 = s_raw + (uint32_t)(min(max(l0 + 1, tlo), thi) * pitch) * 4u;
        const uint32_t p2 = s_raw + (uint32_t)(min(max(l0 + 2, tlo), thi) * pitch) * 4u;
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
            float4 m = v;
            if (c0 + 3 >= N) {  // ragged last group: columns >= N do not take part
              if (c0 + 1 >= N) m.y = -1.f;
              if (c0 + 2 >= N) m.z = -1.f;
              m.w = -1.f;
            }
            const float top = fmaxf(fmaxf(m.x, m.y), fmaxf(m.z, m.w));
            if (top > best_v) {
              best_v = top;
              best_i = r * N + c0 + (m.x == top ? 0 : (m.y == top ? 1 : (m.z == top ? 2 : 3)));
            }
          }
        }
      }
      RSB_STAMP(7);
      float inv = 1.f;
      if (tid == 0) bulk_wait_read();  // the previous image's staged rows have left the stage
      if (!normalize && stage_rows > 0) worker_sync();
      if (normalize) {
        my_best = warp_max_u64(best_v >= 0.f ? best_key(best_v, best_i) : 0ull);
        if (lane == 0) atomicMax(&sm.lbest, my_best);
        worker_sync();  // (G) this band's arg-max complete
        image_best = sm.lbest;
        const unsigned long long theirs = exchange(2, image_best);
        image_best = theirs > image_best ? theirs : image_best;
        inv = 1.f / __uint_as_float((unsigned)(image_best >> 32));
      }
      RSB_STAMP(8);
      if (!f64 && vec) {
        // float32, 16-byte aligned rows: the common case, kept free of the general store's
        // branches (they cost more issue slots than the stores themselves)
        // (running addresses, laundered: the compiler otherwise rebuilds the 64-bit address
        // from the kernel arguments in front of every store)
        float* outf = keep_ptr(reinterpret_cast<float*>(out) + (size_t)(R0 + warp) * opitch + 4 * lane);
        uint32_t s_st = keep_u32(s_stage + (uint32_t)(warp * opitch + 4 * lane) * 4u);
        const uint32_t step = keep_u32((uint32_t)(kWarps * opitch));  // floats between this warp's rows
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
          const int rl = warp + i * kWarps;
          if (rl < H) {
#pragma unroll
            for (int s = 0; s < NS; ++s) {
              if (s >= strips || s * kStrip + 4 * lane >= N) continue;
              float4 v = res[i][s];
              v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
              if (rl < stage_rows) sts_v4(s_st + (uint32_t)(s * kStrip) * 4u, v);
              else *reinterpret_cast<float4*>(outf + s * kStrip) = v;
            }
          }
          outf += step;
          s_st += step * 4u;
        }
      } else {
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
          const int rl = warp + i * kWarps, r = R0 + rl;
          if (rl >= H) continue;
#pragma unroll
          for (int s = 0; s < NS; ++s) {
            const int c0 = s * kStrip + 4 * lane;
            if (s >= strips || c0 >= N) continue;
            float4 v = res[i][s];
            v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
            store_out4(out + (size_t)r * opitch * esz, f64, c0, N, vec, v);
          }
        }
      }
      }  // !kBig
    }
    if (stage_rows > 0) {
      fence_proxy_async();  // my writes to the stage -> visible to the async (TMA) proxy
      bulk_dst = out + (size_t)R0 * opitch * 4;  // issued by thread 0 behind the next barrier
      bulk_bytes = (uint32_t)(stage_rows * opitch * 4);
    }

    RSB_STAMP(9);
    // this image's labels are produced by the label warp during the next image
    if (tid == 0) {
      Fixed2::LabelReq q;
      q.best = image_best; q.img = img; q.pb = pb; q.margin = margin; q.zpad = zpad; q.n = N; q.pad = 0;
      sm.label[buf] = q;
    }
  }
  __syncthreads();
  if (tid == 0 && bulk_dst) {
    bulk_s2g(const_cast<char*>(bulk_dst), stage, bulk_bytes);
    bulk_wait_all();
  }
  if (rank == 0 && warp == kLabelWarp && it > 0) label_job((it - 1) & 1);
  if constexpr (kCluster > 1) cg::this_cluster().sync();  // no CTA exits while its smem is in use
}

// ------------------------------------------------------------------ host side
template <int kCluster, int NS, bool kVar = false, bool kBig = false>
static int launch_v2(const Args& a, size_t smem_bytes, cudaStream_t st) {
  auto k = jet2_kernel<kCluster, NS, kVar, kBig>;
  RSB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  const int sms = sm_count();
  // CTAs are limited to one per SM by registers (512 threads x 128); small images leave
  // shared memory for more, but the register file does not
  const int full = sms / kCluster;
  // a cluster lists its images of the launch in shared memory (kListCap entries): longer batches go in pieces
  const int piece = (kListCap - 2) * full;
  const size_t esz = (a.flags & RSB_JET_OUT_F64) ? 8 : 4;
  for (int i0 = 0; i0 < a.n; i0 += piece) {
    Args b = a;
    b.n = (a.n - i0) < piece ? (a.n - i0) : piece;
    b.params = a.params + (size_t)i0 * a.slots * 6;
    b.zoom_pad = a.zoom_pad + i0;
    b.margin = a.margin + i0;
    b.peak = a.peak + i0;
    if (kVar) {
      b.sizes = a.sizes + i0;
      b.out_offs = a.out_offs + i0;  // (offsets stay relative to a.out)
    } else {
      b.out = reinterpret_cast<char*>(a.out) + (size_t)i0 * a.size * a.size * esz;
    }
    const int nclusters = full < b.n ? full : b.n;
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
    RSB_CUDA(cudaLaunchKernelEx(&cfg, k, b));
    RSB_LAUNCHED();
  }
  return 0;
}

}  // namespace v2

static size_t v2_bytes(int size, int cluster, int stage_rows = 0) {
  const int npad = (size + 3) & ~3;
  const int rows = (size + cluster - 1) / cluster + (cluster > 1 ? 1 : 0);
  return sizeof(v2::Fixed2) + (size_t)npad * 16 + (size_t)(rows + stage_rows) * npad * 4;
}
// rows of a band that are staged for the bulk store: half the band when that fits
static int v2_stage_rows(int size, int cluster, int max_optin) {
  const int half = ((size + cluster - 1) / cluster) / 2;
  return v2_bytes(size, cluster, half) <= (size_t)max_optin ? half : 0;
}

bool v2_supported(int size, int slots, int* cluster, size_t* smem_bytes) {
  int dev = 0, max_optin = 0;
  if (slots > 32) return false;  // per-tile slot sets are 32-bit masks
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess)
    return false;
  if (size <= v2::kMaxBand && v2_bytes(size, 1) <= (size_t)max_optin) {
    *cluster = 1;
    *smem_bytes = v2_bytes(size, 1, v2_stage_rows(size, 1, max_optin));
    return true;
  }
  if (size <= v2::kBigMax && v2_bytes(size, 1) <= (size_t)max_optin) {  // whole image in one CTA, V pass twice
    *cluster = 1;
    *smem_bytes = v2_bytes(size, 1);
    return true;
  }
  if ((size + 1) / 2 <= v2::kMaxBand && v2_bytes(size, 2) <= (size_t)max_optin) {  // (odd sizes: uneven bands)
    *cluster = 2;
    *smem_bytes = v2_bytes(size, 2, v2_stage_rows(size, 2, max_optin));
    return true;
  }
  return false;
}

int v2_launch(const Args& a_in, int cluster, size_t smem_bytes, cudaStream_t st) {
  Args a = a_in;
  a.stage_rows = (int)((smem_bytes - v2_bytes(a.size, cluster)) / ((size_t)((a.size + 3) & ~3) * 4));
#ifdef RSB_TIMELINE
  fprintf(stderr, "v2_launch: size %d cluster %d smem %zu fixed %zu stage_rows %d\n", a.size, cluster, smem_bytes, sizeof(v2::Fixed2), a.stage_rows);
#endif
  const int strips = (a.size + kStrip - 1) / kStrip;  // <= 2 for size <= 256
  if (cluster == 1 && a.size > v2::kMaxBand) return v2::launch_v2<1, 2, false, true>(a, smem_bytes, st);
  if (cluster == 1)
    return strips <= 1 ? v2::launch_v2<1, 1>(a, smem_bytes, st) : v2::launch_v2<1, 2>(a, smem_bytes, st);
  return strips <= 1 ? v2::launch_v2<2, 1>(a, smem_bytes, st) : v2::launch_v2<2, 2>(a, smem_bytes, st);
}

// Variable-size launch: a.size is the largest size of the batch.  One CTA per image when every
// band fits (size <= 128), else a 2-CTA cluster (size <= 256, even or odd).
int v2_launch_var(const Args& a_in, cudaStream_t st) {
  int dev = 0, max_optin = 0;
  RSB_CUDA(cudaGetDevice(&dev));
  RSB_CUDA(cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  Args a = a_in;
  const bool big = a.size > v2::kMaxBand && a.size <= v2::kBigMax && v2_bytes(a.size, 1) <= (size_t)max_optin;
  const int cluster = (a.size <= v2::kMaxBand || big) ? 1 : 2;
  RSB_ARG(a.slots <= v2::kSlots2 && a.size <= 2 * v2::kMaxBand && v2_bytes(a.size, cluster) <= (size_t)max_optin,
          "variable-size jets: sizes up to %d and %d slots only", 2 * v2::kMaxBand, v2::kSlots2);
  const size_t smem = big ? v2_bytes(a.size, 1) : v2_bytes(a.size, cluster, v2_stage_rows(a.size, cluster, max_optin));
  const int npad = (a.size + 3) & ~3;
  a.raw_floats = (int)((smem - sizeof(v2::Fixed2) - (size_t)npad * 16) / 4);
  a.pitch = npad;
  a.stage_rows = 0;
  const int strips = (a.size + kStrip - 1) / kStrip;
  if (big) return v2::launch_v2<1, 2, true, true>(a, smem, st);
  if (cluster == 1)
    return strips <= 1 ? v2::launch_v2<1, 1, true>(a, smem, st) : v2::launch_v2<1, 2, true>(a, smem, st);
  return v2::launch_v2<2, 2, true>(a, smem, st);
}

}  // namespace jet
}  // namespace rsb

#ifdef RSB_TIMELINE
extern "C" int rsb_debug_set_timeline(long long* dev_ptr) {
  RSB_CUDA(cudaMemcpyToSymbol(rsb::jet::v2::g_timeline, &dev_ptr, sizeof(dev_ptr)));
  return 0;
}
#endif
