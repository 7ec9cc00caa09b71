// On-chip jet renderer for sizes 257 .. 512: the jet2.cu design with MORE THAN TWO row bands.
// (reference: src/radiosim/jets/jet.py:144-192, utils/gauss.py:50-97, utils/utils.py:90-177)
//
// One image per thread-block cluster of 8 CTAs; the image's 8-row groups are dealt round-robin, CTA k keeps
// groups k, k + 8, ... in its shared memory (the jet sits around the centre: contiguous bands leave the outer
// CTAs idle).  Splat, margin search, tap table
// and the H pass are band-local exactly as in jet2.cu.  The composite zoom operator only preserves the image
// centre, so with more than two bands an output row's taps can lie in ANOTHER CTA's band: the V pass reads
// them through distributed shared memory (ld.shared::cluster), between two cluster barriers (every band's H
// pass done / every CTA's V reads done).  Margin and arg-max are all-gathered with st.async into per-rank
// slots of every peer (completion on the receiver's mbarrier), as in jet2.cu.
// Replaces, for these sizes, the scratch-image kernel of jet.cu, which streams every pass through HBM.
#include <cooperative_groups.h>

#include "jet_common.cuh"

namespace cg = cooperative_groups;

namespace rsb {
namespace jet {
namespace v3 {

// 16 bytes from ANY CTA's shared memory of the cluster (address from mapa / cluster_map)
__device__ __forceinline__ float4 ldc_v4(uint32_t a) {
  float4 v;
  asm volatile("ld.shared::cluster.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(a));
  return v;
}

#define RSB_STAMP(k) do {} while (0)
#define RSB_STAMP_T(k, t) do {} while (0)
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
constexpr int kMaxBand = 64;    // rows per CTA band (V-pass results live in registers: 4 rows per warp x 4 strips)
constexpr int kRowsPerWarp = kMaxBand / kWarps;
constexpr int kMaxGroups = kMaxBand / kTileRows;
constexpr int kMaxCand = 64;
constexpr int kMaxTiles = kMaxGroups * 8;  // 64-column tiles of a band (size <= 512)
constexpr int kMaxRec = 8;                 // tiles per warp and image (work list length)
constexpr unsigned kRecEnd = 0xffffu;      // list terminator in rec.y

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
  unsigned long long xval[3][2][8]; // every CTA's margin / (unused) / arg-max key, [kind][use parity][rank]
  uint64_t xbar[3][2];              // mbarriers the peers' st.async complete: margin, (unused), best; [use parity]
  unsigned long long lbest;         // this CTA's running arg-max key
  int lmargin;                      // this CTA's running margin
  int nlive[2];
  int next_row;
  int foot_lo, foot_hi;             // output coordinates with a non-zero tap
  float delta[2];
  int dlo[2];                       // no pixel closer than this to the border exceeds 0.01
  struct LabelReq { unsigned long long best; int img, pb, margin, zpad, n, pad; } label[2];  // [image parity]
  alignas(16) float geo[kSlots2 * 16];  // build_coefs2 scratch (SlotGeo per live slot, read as float4)
  int ncand;                        // borderline pixels awaiting a float64 verdict
  int cand[kMaxCand];               // r << 16 | c
  alignas(16) uint2 rec[2][kWarps][kMaxRec];    // per-warp work lists: {slots that touch the tile, row group << 8 | column tile}
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

#define RSB_BSTAMP(k) do {} while (0)
__device__ inline void build_coefs2(Fixed2& sm, int cb, const double* P, int slots, int N, int R0,
                                    int H, int lane, bool zoom, int first_worker, int bands, int rank,
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
        // (8-row groups are dealt round-robin to the CTAs of the cluster: local group g = absolute g * bands + rank)
        const int r0 = (g * bands + rank) * kTileRows, tc0 = tc * kTileCols;
        const int rend = min(N, r0 + kTileRows);
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

template <int kCluster, int NS>
__global__ void __launch_bounds__(kThreads, 1) jet3_kernel(const Args a) {
  static_assert(kCluster >= 2 && kCluster <= 8 && NS <= 4, "band kernel: 2..8 CTAs per image, up to 512 columns");
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Fixed2& sm = *reinterpret_cast<Fixed2*>(smem_raw);
  const int slots = a.slots;
  const int npad_max = (a.size + 3) & ~3;
  // composite taps, one record per output coordinate: {w0, w1, w2, first tap (absolute, as int bits)}
  float4* tab = reinterpret_cast<float4*>(smem_raw + sizeof(Fixed2));  // [npad_max]
  float* raw = reinterpret_cast<float*>(tab + npad_max);               // [Hb][pitch]: rows R0 .. R0 + H - 1

  const uint32_t s_raw = keep_u32(smem_u32(raw)), s_tab = keep_u32(smem_u32(tab));
  const int tid = keep_s32((int)threadIdx.x), lane = tid & 31, warp = tid >> 5;
  const int rank = keep_s32((kCluster > 1) ? (int)(blockIdx.x % kCluster) : 0);
  const int cid = blockIdx.x / kCluster, nclusters = gridDim.x / kCluster;
  const bool f64 = (a.flags & RSB_JET_OUT_F64) != 0;
  const bool normalize = (a.flags & RSB_JET_NORMALIZE) != 0;
  const size_t esz = f64 ? 8 : 4;
  // ---- geometry of the image (fixed for the whole launch) ----
  int N, npad, pitch, opitch;  // size, size rounded up to 4, raw row pitch, output row pitch (elements)
  int Hb, H, R0, base;         // rows per band, rows of THIS band, its first absolute row (= base: raw row 0)
  int strips, tcols, stage_rows;
  bool vec;
  float* stage;
  uint32_t s_stage;
  auto geometry = [&](int n_) {
    N = n_;
    npad = (N + 3) & ~3;
    pitch = a.pitch;
    opitch = N;
    // 8-row groups dealt round-robin: CTA k holds groups k, k + kCluster, ... (the jet sits around the image
    // centre: contiguous bands would leave the outer CTAs idle).  Local row lr <-> absolute row
    // ((lr >> 3) kCluster + k) 8 + (lr & 7).
    {
      const int G = (N + kTileRows - 1) / kTileRows;          // groups of the image
      const int ngl = (G - rank + kCluster - 1) / kCluster;   // ... of this CTA (>= 1 for N >= 8 kCluster)
      const int last = ((ngl - 1) * kCluster + rank) * kTileRows;
      Hb = ((G + kCluster - 1) / kCluster) * kTileRows;       // rows allocated per CTA
      H = (ngl - 1) * kTileRows + min(kTileRows, N - last);   // local rows in use
    }
    R0 = 0;      // (local row 0; absolute rows come from abs_row())
    base = 0;
    strips = (N + kStrip - 1) / kStrip;           // 128-column strips (H/V pass)
    tcols = (N + kTileCols - 1) / kTileCols;      // 64-column tiles (phase 1)
    vec = (opitch & 3) == 0 && (reinterpret_cast<uintptr_t>(a.out) & 15) == 0;
    // float32 output with contiguous rows: the first stage_rows rows of the finished band are staged in
    // shared memory and leave through one bulk async copy that drains while the next image is rendered
    stage_rows = 0;  // (a CTA's rows are not contiguous in the output: no staged bulk store)
    stage = raw + (size_t)Hb * pitch;
    s_stage = keep_u32(smem_u32(stage));
  };
  geometry(a.size);
  auto abs_row = [&](int lr) -> int { return (((lr >> 3) * kCluster + rank) << 3) + (lr & 7); };
  int xuse[3] = {0, 0, 0};  // uses so far of each exchange mbarrier (same in every CTA)
  // all-gather of one 64-bit value per CTA (all threads call; the value is CTA-uniform): thread 0 writes it
  // into slot [rank] of every peer with st.async, completion counted on the receiver's mbarrier; returns
  // the minimum (kind 0: margins) or the maximum (kind 2: arg-max keys) over the cluster
  auto exchange = [&](int kind, unsigned long long mine) -> unsigned long long {
    const int par = xuse[kind] & 1;
    // (two mbarriers per kind, by use parity: a CTA that is one exchange ahead completes bytes on the OTHER
    // barrier, never on the one a slower CTA is still waiting on; nobody can be two ahead)
    if (tid == 0) {
      mbar_expect_tx(&sm.xbar[kind][par], 8u * (kCluster - 1));
      const uint32_t slot = smem_u32(&sm.xval[kind][par][rank]), bar = smem_u32(&sm.xbar[kind][par]);
#pragma unroll
      for (int p = 0; p < kCluster; ++p)
        if (p != rank) st_async_b64(cluster_map(slot, (uint32_t)p), mine, cluster_map(bar, (uint32_t)p));
    }
    mbar_wait(&sm.xbar[kind][par], (uint32_t)((xuse[kind] >> 1) & 1));
    xuse[kind]++;
    unsigned long long r = mine;
#pragma unroll
    for (int p = 0; p < kCluster; ++p) {
      if (p == rank) continue;
      const unsigned long long v = sm.xval[kind][par][p];
      r = (kind == 0) ? (v < r ? v : r) : (v > r ? v : r);
    }
    return r;
  };

  const uint32_t pbytes = (uint32_t)slots * RSB_JET_PARAM_STRIDE * 8u;
  const char* bulk_dst = nullptr;  // (thread 0) where the staged rows of the previous image go
  uint32_t bulk_bytes = 0;         // ... and how many bytes of the stage they are

  if (tid == 0) {
    mbar_init(&sm.bar[0], 1);
    mbar_init(&sm.bar[1], 1);
    mbar_init(&sm.bar[2], 1);
    for (int k = 0; k < 3; ++k) {
      mbar_init(&sm.xbar[k][0], 1);
      mbar_init(&sm.xbar[k][1], 1);
    }
    fence_mbar_init();
    sm.lmargin = N;
    sm.lbest = 0ull;
    sm.next_row = 0;
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
  // warp 1 only has its side job (the labels) in the cluster's first CTA; elsewhere it renders
  const int first_worker = (rank == 0) ? 2 : 1;
  if (warp == 0 && cid < a.n) {  // prologue: coefficients of the first image
    mbar_wait(&sm.bar[0], 0);
    build_coefs2(sm, 0, sm.params[0], slots, N, R0, H, lane, a.zoom_pad[cid] >= 0, first_worker, kCluster, rank);
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
      peak = peak64(sm.pix64, sm.params[q.pb], slots, q.n, q.zpad >= 0, q.zpad, q.margin, q.best, lane);
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
    __syncthreads();  // (A) coef/span/nlive of this image visible; raw free for reuse; stage complete
    if (tid == 0 && bulk_dst) {
      bulk_s2g(const_cast<char*>(bulk_dst), stage, bulk_bytes);
      bulk_dst = nullptr;
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
        if (img + nclusters < a.n) {
          const int nb = (it + 1) % 3;
          mbar_wait(&sm.bar[nb], ((it + 1) / 3) & 1);
          build_coefs2(sm, buf ^ 1, sm.params[nb], slots, N, R0, H, lane, a.zoom_pad[img + nclusters] >= 0,
                       first_worker, kCluster, rank);
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
#define DBG_LAP(acc) do {} while (0)
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
      const int r0 = (g * kCluster + rank) * kTileRows, tc0 = tc * kTileCols, c0 = tc0 + 2 * lane;
      const uint32_t s_dst = s_raw + (uint32_t)(g * kTileRows * pitch + c0) * 4u;  // local rows g 8 ..
      const int rend = min(N, r0 + kTileRows);  // rows of this tile that exist
      const float r0f = (float)r0, c0f = (float)c0;
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

    RSB_STAMP(1);  // this warp is out of tiles
    int margin = 0;
    char* out = reinterpret_cast<char*>(a.out) +
                (size_t)img * N * N * esz;
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
      for (int e = tid; e < H * q4; e += kWorkers) {
        const int rl = e / q4, c0 = (e - rl * q4) << 2;
        const int r = abs_row(rl);
        float4 v = lds_v4(s_raw + (uint32_t)(rl * pitch + c0) * 4u);
        v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
        if (!f64 && vec) *reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + (size_t)r * opitch + c0) = v;
        else store_out4(out + (size_t)r * opitch * esz, f64, c0, N, vec, v);
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


      // H pass, in place: row r of raw becomes row r of T.  Rows are dealt statically; lane <-> ADJACENT
      // columns (c = strip + 32 j + lane): a warp's gathers touch consecutive words (composite stride <= 2).
      // With up to four strips the tap records (16 per lane) do not stay in registers: they come from the
      // table again for every row.
      {
        bool act[NS];
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          const int sc0 = s * kStrip;
          act[s] = (s < strips) && sc0 <= fhi && sc0 + kStrip - 1 >= flo;
        }
        for (int lr = warp; lr < H; lr += kWarps) {
          const int r = abs_row(lr);
          if (r < margin || r >= N - margin) continue;  // (warp-uniform) only the crop's rows are resampled
          const uint32_t s_row = s_raw + (uint32_t)(lr * pitch) * 4u;
          float res[NS][4];
          // two strips at a time: their eight tap records first, then the 24 gathers, then the arithmetic --
          // independent loads in flight instead of sixteen dependent record -> gather -> fma chains
#pragma unroll
          for (int s0 = 0; s0 < NS; s0 += 2) {
            float4 tp[2][4];
#pragma unroll
            for (int q = 0; q < 2; ++q)
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const int cj = (s0 + q) * kStrip + 32 * j + lane;
                tp[q][j] = (s0 + q < NS) ? lds_v4(s_tab + (uint32_t)min(cj, npad - 1) * 16u) : make_float4(0.f, 0.f, 0.f, 0.f);
              }
            float v[2][4][3];
#pragma unroll
            for (int q = 0; q < 2; ++q)
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const uint32_t src = s_row + (uint32_t)__float_as_int(tp[q][j].w) * 4u;
                v[q][j][0] = lds_f32(src);
                v[q][j][1] = lds_f32(src + 4u);
                v[q][j][2] = lds_f32(src + 8u);
              }
#pragma unroll
            for (int q = 0; q < 2; ++q)
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (s0 + q < NS)
                  res[s0 + q][j] = fmaf(tp[q][j].z, v[q][j][2], fmaf(tp[q][j].y, v[q][j][1], tp[q][j].x * v[q][j][0]));
          }
          __syncwarp();
#pragma unroll
          for (int s = 0; s < NS; ++s) {
            if (act[s]) {
              const uint32_t s_out = s_row + (uint32_t)(s * kStrip + lane) * 4u;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (s * kStrip + 32 * j + lane < N) sts_f32(s_out + 128u * j, res[s][j]);
            }
          }
        }
      }
      RSB_STAMP(4);
      worker_sync();  // (F) T complete in this band
      RSB_STAMP(5);
      cg::this_cluster().sync();  // every band's T is complete: the V pass may read any CTA's rows

      RSB_STAMP(6);
      const int vr0 = flo, vr1 = fhi;  // inclusive footprint rows
      // absolute rows that hold T: the crop [margin, N - margin) -- in whichever CTA's band they lie
      const int tlo = margin, thi = max(N - margin - 1, margin);
      const uint32_t raw_off = s_raw;  // same offset in every CTA's shared window
      // address of T[row][c0 ..] in the shared memory of the CTA that owns the row
      auto t_row = [&](int row) -> uint32_t {
        const int g = row >> 3, owner = g % kCluster, lr = ((g / kCluster) << 3) + (row & 7);
        return cluster_map(raw_off + (uint32_t)(lr * pitch) * 4u, (uint32_t)owner);
      };
      // V pass: results stay in registers (4 rows x NS strips of 4 columns per lane).  A warp owns four
      // CONSECUTIVE output rows (half of one of this CTA's 8-row groups): their tap rows overlap, so the
      // union of them -- 3 k + 3 rows for a composite stride k, against 12 -- is fetched once through
      // distributed shared memory and every fetched row is accumulated into the output rows that use it.
      // Taps are visited in increasing row order, i.e. in each output's own tap order: same sums as
      // fma(w2, t2, fma(w1, t1, w0 t0)).
      float4 res[kRowsPerWarp][NS];
      // running maximum of this lane: value and pixel index (its pixels come in increasing
      // index order, so '>' keeps the first of equal values)
      best_v = -1.f;
      best_i = 0;
      auto my_row = [&](int i) -> int { return ((warp >> 1) << 3) + ((warp & 1) << 2) + i; };  // local row
      {
        int tb[kRowsPerWarp];            // first tap row of each output row (absolute)
        float tw[kRowsPerWarp][3];       // its three weights (all zero: row off the footprint)
        int umin = 0x7fffffff, umax = -1;
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
          const int lr = my_row(i), r = abs_row(lr);
#pragma unroll
          for (int s = 0; s < NS; ++s) res[i][s] = make_float4(0.f, 0.f, 0.f, 0.f);
          tb[i] = 0;
          tw[i][0] = tw[i][1] = tw[i][2] = 0.f;
          if (lr < H && r >= vr0 && r <= vr1) {  // warp-uniform
            const float4 tp = lds_v4(s_tab + (uint32_t)r * 16u);
            tb[i] = __float_as_int(tp.w);
            tw[i][0] = tp.x; tw[i][1] = tp.y; tw[i][2] = tp.z;
            // (taps with a non-zero weight lie in the crop [tlo, thi]: only those are fetched)
#pragma unroll
            for (int t = 0; t < 3; ++t)
              if (tw[i][t] != 0.f) {
                umin = min(umin, tb[i] + t);
                umax = max(umax, tb[i] + t);
              }
          }
        }
        umin = max(umin, tlo);
        umax = min(umax, thi);
        for (int u = umin; u <= umax; ++u) {
          float wu[kRowsPerWarp];
          bool any = false;
#pragma unroll
          for (int i = 0; i < kRowsPerWarp; ++i) {
            const int o = u - tb[i];
            wu[i] = (o == 0) ? tw[i][0] : (o == 1) ? tw[i][1] : (o == 2) ? tw[i][2] : 0.f;
            any = any || wu[i] != 0.f;
          }
          if (!any) continue;  // (a gap between two output rows' taps)
          const uint32_t pu = t_row(u);
#pragma unroll
          for (int s = 0; s < NS; ++s) {
            const int sc0 = s * kStrip, c0 = sc0 + 4 * lane;
            if (!(s < strips && sc0 <= fhi && sc0 + kStrip - 1 >= flo) || c0 >= N) continue;
            const float4 t = ldc_v4(pu + (uint32_t)c0 * 4u);
            const f2 tlo2 = f2_pack(t.x, t.y), thi2 = f2_pack(t.z, t.w);
#pragma unroll
            for (int i = 0; i < kRowsPerWarp; ++i) {
              const f2 W = f2_pack(wu[i], wu[i]);
              float4& v = res[i][s];
              f2_unpack(f2_fma(W, tlo2, f2_pack(v.x, v.y)), v.x, v.y);
              f2_unpack(f2_fma(W, thi2, f2_pack(v.z, v.w)), v.z, v.w);
            }
          }
        }
        if (normalize) {
#pragma unroll
          for (int i = 0; i < kRowsPerWarp; ++i) {
            const int lr = my_row(i), r = abs_row(lr);
            if (lr >= H || r < vr0 || r > vr1) continue;
#pragma unroll
            for (int s = 0; s < NS; ++s) {
              const int sc0 = s * kStrip, c0 = sc0 + 4 * lane;
              if (!(s < strips && sc0 <= fhi && sc0 + kStrip - 1 >= flo) || c0 >= N) continue;
              float4 m = res[i][s];
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
      }
      RSB_STAMP(7);
      cg::this_cluster().sync();  // every CTA's V reads are done: the bands may be rendered into again
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
        // float32, 16-byte aligned rows: the common case, kept free of the general store's branches
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
          const int rl = my_row(i);
          if (rl < H) {
            float* outf = reinterpret_cast<float*>(out) + (size_t)abs_row(rl) * opitch + 4 * lane;
#pragma unroll
            for (int s = 0; s < NS; ++s) {
              if (s >= strips || s * kStrip + 4 * lane >= N) continue;
              float4 v = res[i][s];
              v.x *= inv; v.y *= inv; v.z *= inv; v.w *= inv;
              *reinterpret_cast<float4*>(outf + s * kStrip) = v;
            }
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < kRowsPerWarp; ++i) {
          const int rl = my_row(i);
          if (rl >= H) continue;
          const int r = abs_row(rl);
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
template <int kCluster, int NS>
static int launch_v3(const Args& a, size_t smem_bytes, cudaStream_t st) {
  auto k = jet3_kernel<kCluster, NS>;
  RSB_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes));
  cudaLaunchConfig_t cfg = {};
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
  // clusters that can be resident at once (a cluster lives inside one GPC): the kernel is persistent
  cfg.gridDim = dim3((unsigned)(sm_count() / kCluster * kCluster));
  int nclusters = 0;
  RSB_CUDA(cudaOccupancyMaxActiveClusters(&nclusters, k, &cfg));
  if (nclusters < 1) nclusters = 1;
  if (nclusters > a.n) nclusters = a.n;
  cfg.gridDim = dim3((unsigned)(nclusters * kCluster));
  RSB_CUDA(cudaLaunchKernelEx(&cfg, k, a));
  RSB_LAUNCHED();
  return 0;
}

}  // namespace v3

constexpr int kV3Cluster = 8;
static size_t v3_bytes(int size, int stage_rows) {
  const int npad = (size + 3) & ~3;
  const int groups = (size + 7) / 8;
  const int rows = ((groups + kV3Cluster - 1) / kV3Cluster) * 8;  // 8-row groups dealt round-robin
  return sizeof(v3::Fixed2) + (size_t)npad * 16 + (size_t)(rows + stage_rows) * npad * 4;
}

// sizes 401 .. 512 (at most 64 rows per CTA), at most 32 slots.  Below ~400 pixels the scratch-image kernel of
// jet.cu is faster (measured crossover: 384^2 1.01 vs 1.06 M images/s, 448^2 0.80 vs 0.70): the per-image cost of
// eight CTAs' barriers and of the V pass's distributed-shared-memory reads does not shrink with the image.
bool v3_supported(int size, int slots, size_t* smem_bytes) {
  int dev = 0, max_optin = 0;
  if (slots > 32 || size <= 400 || size > kV3Cluster * v3::kMaxBand) return false;
  if (cudaGetDevice(&dev) != cudaSuccess) return false;
  if (cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev) != cudaSuccess)
    return false;
  if (v3_bytes(size, 0) > (size_t)max_optin) return false;
  *smem_bytes = v3_bytes(size, 0);  // (no staged bulk store: a CTA's rows are scattered over the image)
  return true;
}

int v3_launch(const Args& a_in, size_t smem_bytes, cudaStream_t st) {
  Args a = a_in;
  a.stage_rows = (int)((smem_bytes - v3_bytes(a.size, 0)) / ((size_t)((a.size + 3) & ~3) * 4));
  return v3::launch_v3<kV3Cluster, 4>(a, smem_bytes, st);  // four 128-column strips for 401 .. 512
}

}  // namespace jet
}  // namespace rsb
