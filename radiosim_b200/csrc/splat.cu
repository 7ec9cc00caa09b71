// Tiled Gaussian splat for survey images (many components on large grids), one kernel from
// parameters to the five normalised output planes.
//
// Reference: src/radiosim/survey/survey.py:49-122 -- every source is evaluated on the full
// N x N grid (gauss.py:50-97) and added to its class plane; the image is the plane sum
// divided by its maximum, the last plane is 1 - image.
//
// Here:
//   prep_kernel   : float64 source parameters -> 48-byte float32 conic records + the bounding
//                   box of the > tau footprint (one thread per source).
//   tile_kernel   : persistent, one CTA per SM, 64 x 128 pixel tiles handed out by a ticket
//                   counter.  Per tile:
//                     * one warp builds -- one tile AHEAD, while the others render -- the list
//                       of sources and sub-jets whose box meets the tile (ordered: the sums
//                       do not depend on scheduling);
//                     * their records are gathered into shared memory;
//                     * 96 warp tasks (class x 4-row x 64-column strip) cull the staged list
//                       with an exact ellipse-vs-rectangle test (one ballot per 32 sources) and
//                       accumulate 8 pixels per lane in registers: rows in blocks of four from
//                       two exponentials per column (half the MUFU work of one exp per pixel),
//                       packed FP32x2 arithmetic; sub-jets (class 0, rendered beforehand by the
//                       jet kernel) are pasted with the reference's clipping;
//                     * the three class planes of the tile stay in shared memory; the tile's
//                       maximum of the plane sum goes to a per-image atomicMax and the tile
//                       "arrives" at a per-image counter;
//                     * the tile is FLUSHED one tile later, when every tile of its image has
//                       arrived (two tile buffers: the wait hides behind the next tile): the
//                       five planes x, y[0..3] are scaled by 1/max and written once with 128-bit
//                       stores.
// No unnormalised plane ever goes to global memory: DRAM traffic = the 5 output planes.
#include "jet_common.cuh"

namespace rsb {
namespace splat {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kTileRows = 64, kTileCols = 128;   // CTA tile
constexpr int kSubRows = 8, kSubCols = 64;       // warp task: 8 rows x 64 columns, 2 columns per lane
constexpr int kGroups = kTileRows / kSubRows;    // 8 row groups
constexpr int kHalves = kTileCols / kSubCols;    // 2
constexpr int kSub = kGroups * kHalves;          // 16 sub-tiles = one per warp
constexpr int kSrcTasks = 2 * kSub;              // class (wide, narrow) x sub-tile
constexpr int kRectTasks = kSub;
static_assert(kSub == kWarps, "one wide task per warp");
constexpr int kRecCap = 512;                     // staged source records per round (24 KB)
constexpr int kRectCap = 128;                    // staged sub-jet rectangles per round
constexpr int kTilePx = kTileRows * kTileCols;
constexpr int kBuilderWarp = kWarps - 1;

// One source as the tile loop reads it (warp-uniform):
//   g(r,c) = 2^(la + A dc^2 + B dc dr + C dr^2),  dc = c - (icf + fc), dr = r - (irf + fr)
// Rows come in blocks of four from two exponentials per column,
//   E(i) = E0 U^i 2^(C (i^2 - i)),  E0 = g(r_b, c),  U = 2^(B dc + 2 C dr_b + C),  k2 = 2^(2C).
// Sources this does not suit (sigma below ~0.6 pixel, non-finite numbers) carry k2 < 0 and
// la = the amplitude itself: one exponential per pixel.
// Culling: the source touches a rectangle iff max over it of q = A dc^2 + B dc dr + C dr^2
// reaches T2s = log2(tau / amp); g1 = B / 2C, g2 = B / 2A locate the maximum along an edge.
struct __align__(16) Rec {
  float A, B, C, la;
  float irf, fr, icf, fc;
  float k2, T2s, g1, g2;
};
static_assert(sizeof(Rec) == 48, "Rec layout");

struct JetRect {  // one pasted sub-jet (survey.py:56-76)
  long long pix;  // offset (floats) of its first pixel in the packed jet buffer
  int size, row0, col0, pitch;  // edge, image position of its pixel (0,0), floats between its rows
};
static_assert(sizeof(JetRect) == 24, "JetRect layout");

// ---- prep: parameters -> records + bounding boxes -------------------------------------------
__global__ void prep_kernel(const double* __restrict__ params, int total, int N, double tau,
                            Rec* __restrict__ rec, int4* __restrict__ bbox) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= total) return;
  const double* p = params + (size_t)k * 6;
  const double amp = p[0], x = p[1], y = p[2], sx = p[3], sy = p[4];
  Rec c;
  c.A = c.B = c.C = 0.f; c.la = 0.f; c.irf = c.fr = c.icf = c.fc = 0.f;
  c.k2 = -1.f; c.T2s = 0.f; c.g1 = c.g2 = 0.f;
  int4 bb = make_int4(0x7fffffff, -1, 0x7fffffff, -1);  // r0 r1 c0 c1, empty: never listed
  if (amp != 0.0 && fabs(amp) > tau) {
    double sr, cr;
    sincos(p[5], &sr, &cr);
    const double iu = 1.0 / (sy * sy), iv = 1.0 / (sx * sx);
    const double Mrr = cr * cr * iu + sr * sr * iv, Mcc = sr * sr * iu + cr * cr * iv;
    const double Mrc = sr * cr * (iv - iu);
    const double h = 0.5 * 1.4426950408889634;
    c.A = (float)(-h * Mcc);
    c.C = (float)(-h * Mrr);
    c.B = (float)(-2.0 * h * Mrc);
    const bool blocks = amp > 0.0 && c.C >= -2.f && isfinite(c.A) && isfinite(c.B) && isfinite(c.C);
    c.la = blocks ? (float)log2(amp) : (float)amp;
    c.k2 = blocks ? exp2f(2.f * c.C) : -1.f;
    c.T2s = (float)log2(tau / fabs(amp));
    c.g1 = (float)(Mrc / Mrr);
    c.g2 = (float)(Mrc / Mcc);
    const double yr = rint(y), xr = rint(x);
    c.irf = (float)yr; c.fr = (float)(y - yr);
    c.icf = (float)xr; c.fc = (float)(x - xr);
    const double T2 = 2.0 * log(fabs(amp) / tau), det = Mrr * Mcc - Mrc * Mrc;
    const double er = sqrt(T2 * Mcc / det), ec = sqrt(T2 * Mrr / det);
    if (isfinite(er) && isfinite(ec)) {
      const double lo_r = fmax(floor(y - er) - 1.0, 0.0), hi_r = fmin(ceil(y + er) + 1.0, N - 1.0);
      const double lo_c = fmax(floor(x - ec) - 1.0, 0.0), hi_c = fmin(ceil(x + ec) + 1.0, N - 1.0);
      if (lo_r <= hi_r && lo_c <= hi_c) bb = make_int4((int)lo_r, (int)hi_r, (int)lo_c, (int)hi_c);
    } else {  // degenerate widths: never cull
      bb = make_int4(0, N - 1, 0, N - 1);
      c.T2s = -INFINITY;  // every rectangle test says "touches"
    }
  }
  rec[k] = c;
  bbox[k] = bb;
}

struct TileArgs {
  const Rec* rec;            // [total] records, grouped per image: class 1 (wide) then class 2 (narrow)
  const int4* bbox;          // [total]
  const int32_t* offs;       // [n][3]: first class-1 record, first class-2 record, end
  const JetRect* jets;       // [total jets]
  const int32_t* jet_offs;   // [n + 1]
  const float* jet_pix;      // packed sub-jet images (already normalised)
  unsigned* peak;            // [n] float bits of max(sum of planes), zero-initialised
  unsigned* arrived;         // [n] tiles of the image whose maximum is in `peak`, zero-initialised
  unsigned* ticket;          // [1] next tile, zero-initialised
  const int* order;          // [n][tiles] ticket -> tile of its image, most expensive tiles first
  void* out_x;               // [n][N][N]
  void* out_y;               // [n][4][N][N]
  int n, N, tiles_x, tiles;  // tiles per image row / per image
  int direct;                // images with more tiles than the grid has CTAs cannot be held on chip
                             // until their maximum is known: tiles are written unnormalised at
                             // once (background plane = image) and norm_inplace_kernel finishes
};

struct ListInfo {            // one round of one tile, as the builder leaves it
  int ticket;                // tile (image * tiles + tile); >= n * tiles: no more work
  int nsrc, nwide, nrect;    // staged sources (the first nwide are class 1), staged rectangles
  int src_resume, rect_resume;  // where the next round continues; -1: this list is complete
  int pad[2];
};

struct __align__(16) Smem {
  float plane[2][3][kTilePx];         // two tile buffers x (jets, wide, narrow)
  Rec rec[kRecCap];
  JetRect rect[kRectCap];
  unsigned list[2][kRecCap];          // source indices of the current / the next tile
  unsigned rlist[2][kRectCap];
  ListInfo info[2];
  float red[kWarps];
  float inv;
  int task;
};

// index of the lowest set bit on the FMA pipe (FLO/POPC queue behind the exponentials)
__device__ __forceinline__ int lowest_bit(unsigned x) {
  return (int)(__float_as_uint(__uint2float_rn(x & (0u - x))) >> 23) - 127;
}

// One warp: the sources [s_begin, o3) and rectangles [j_begin, j1) whose box meets the tile,
// in index order, at most kRecCap / kRectCap of them (whole groups of 32 only).
__device__ inline void build_round(const TileArgs& a, int ticket, int s_begin, int o2, int o3,
                                   int j_begin, int j1, int R0, int C0, unsigned* list,
                                   unsigned* rlist, ListInfo* info, int lane) {
  const int R1 = R0 + kTileRows - 1, C1 = C0 + kTileCols - 1;
  int ns = 0, nw = 0, resume = -1;
  const unsigned lt = (1u << lane) - 1u;
  for (int s = s_begin; s < o3 && resume < 0; s += 128) {
    int4 bb[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {  // four independent loads in flight
      const int idx = s + 32 * u + lane;
      bb[u] = idx < o3 ? a.bbox[idx] : make_int4(0x7fffffff, -1, 0x7fffffff, -1);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int idx = s + 32 * u + lane;
      const bool hit = !(bb[u].x > R1 || bb[u].y < R0 || bb[u].z > C1 || bb[u].w < C0);
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (resume >= 0 || m == 0u) continue;
      const int cnt = __popc(m);
      if (ns + cnt > kRecCap) { resume = s + 32 * u; continue; }
      if (hit) list[ns + __popc(m & lt)] = (unsigned)idx;
      nw += __popc(__ballot_sync(0xffffffffu, hit && idx < o2));
      ns += cnt;
    }
  }
  int nr = 0, rresume = -1;
  for (int j = j_begin; j < j1 && rresume < 0; j += 32) {
    const int idx = j + lane;
    bool hit = false;
    if (idx < j1) {
      const JetRect jr = a.jets[idx];
      hit = !(jr.row0 > R1 || jr.row0 + jr.size <= R0 || jr.col0 > C1 || jr.col0 + jr.size <= C0);
    }
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    const int cnt = __popc(m);
    if (nr + cnt > kRectCap) { rresume = j; break; }
    if (hit) rlist[nr + __popc(m & lt)] = (unsigned)idx;
    nr += cnt;
  }
  __syncwarp();
  if (lane == 0) {
    ListInfo li;
    li.ticket = ticket; li.nsrc = ns; li.nwide = nw; li.nrect = nr;
    li.src_resume = resume; li.rect_resume = rresume; li.pad[0] = li.pad[1] = 0;
    *info = li;
  }
}

// Take the next ticket and build the first round of its list (one warp).
// A CTA holds a finished tile (waiting for its image's maximum) while it renders the next one.
// It must never hold, at the same time, an UNRENDERED ticket of the image that finished tile
// waits for -- it would wait for itself.  So a ticket of image `avoid_img` is not taken here
// (kDeferred); the caller takes it once the finished tile has been flushed.  With that rule and
// tiles per image <= CTAs of the grid no set of CTAs can wait on each other: the lowest
// incomplete image's unrendered tickets are held by CTAs whose finished tile lies in a lower,
// complete image, and if some of its tickets are still untaken, not every CTA can hold two of
// its tiles.
constexpr int kDeferred = -2;
__device__ inline void build_next(const TileArgs& a, unsigned* list, unsigned* rlist, ListInfo* info,
                                  int lane, int avoid_img) {
  const int total = a.n * a.tiles;
  int t = 0;
  if (lane == 0) {
    // The counter only grows and avoid_img is the image of a ticket this CTA already holds, so
    // once the counter has left that image no later ticket can fall into it: a plain read
    // decides, then atomicAdd takes whatever ticket is next.
    bool defer = false;
    if (avoid_img >= 0) {
      const unsigned seen = *(volatile unsigned*)a.ticket;
      defer = (int)seen < total && (int)(seen / (unsigned)a.tiles) == avoid_img;
    }
    t = defer ? kDeferred : (int)atomicAdd(a.ticket, 1u);
    if (t > total) t = total;
  }
  t = __shfl_sync(0xffffffffu, t, 0);
  if (t >= total || t == kDeferred) {
    if (lane == 0) {
      ListInfo li;
      li.ticket = t; li.nsrc = li.nwide = li.nrect = 0; li.src_resume = li.rect_resume = -1;
      li.pad[0] = li.pad[1] = 0;
      *info = li;
    }
    return;
  }
  const int img = t / a.tiles, tile = a.order[t];
  const int ty = tile / a.tiles_x, tx = tile - ty * a.tiles_x;
  build_round(a, t, a.offs[img * 3], a.offs[img * 3 + 1], a.offs[img * 3 + 2], a.jet_offs[img],
              a.jet_offs[img + 1], ty * kTileRows, tx * kTileCols, list, rlist, info, lane);
}

// ---- one source over this lane's 8 x 2 pixels ---------------------------------------------
__device__ __forceinline__ void source_math(const float4& cq, const float4& cc, const float2& ck,
                                            float r0f, float c0f, f2 (&acc)[kSubRows]) {
  const float d0 = (c0f - cc.z) - cc.w;
  const f2 d = f2_pack(d0, d0 + 1.f);
  const f2 Q = f2_mul(f2_pack(cq.y, cq.y), d);
  const float dr0 = (r0f - cc.x) - cc.y;
  if (ck.x > 0.f) {
    // rows in two blocks of four, two exponentials per column and block (see Rec)
    const f2 Pq = f2_fma(f2_mul(f2_pack(cq.x, cq.x), d), d, f2_pack(cq.w, cq.w));  // A dc^2 + la
    const float k3s = ck.x * ck.x * ck.x;
    const f2 k2 = f2_pack(ck.x, ck.x), k3 = f2_pack(k3s, k3s);
#pragma unroll
    for (int b = 0; b < kSubRows; b += 4) {
      const float dr = dr0 + (float)b;
      const float Cq = cq.z * dr * dr;
      const float tw = fmaf(2.f * cq.z, dr, cq.z);  // 2 C dr + C
      float ta, tb, wa, wb;
      f2_unpack(f2_fma(f2_pack(dr, dr), Q, f2_add(Pq, f2_pack(Cq, Cq))), ta, tb);
      f2_unpack(f2_add(Q, f2_pack(tw, tw)), wa, wb);
      // (the clamp never binds where the block holds a pixel above the cut-off; it keeps U^3 finite)
      const f2 E0 = f2_pack(ex2(ta), ex2(tb));
      const f2 U = f2_pack(ex2(fminf(wa, 40.f)), ex2(fminf(wb, 40.f)));
      const f2 U2 = f2_mul(U, U);
      acc[b] = f2_add(acc[b], E0);
      acc[b + 1] = f2_fma(E0, U, acc[b + 1]);
      acc[b + 2] = f2_fma(E0, f2_mul(U2, k2), acc[b + 2]);
      acc[b + 3] = f2_fma(E0, f2_mul(f2_mul(U2, U), k3), acc[b + 3]);
    }
  } else {
    // one exponential per pixel; cq.w holds the amplitude itself
    const f2 Pq = f2_mul(f2_mul(f2_pack(cq.x, cq.x), d), d);
    const f2 amp2 = f2_pack(cq.w, cq.w);
#pragma unroll
    for (int i = 0; i < kSubRows; ++i) {
      const float dri = dr0 + (float)i;
      const float Cq = cq.z * dri * dri;
      float ta, tb;
      f2_unpack(f2_fma(f2_pack(dri, dri), Q, f2_add(Pq, f2_pack(Cq, Cq))), ta, tb);
      acc[i] = f2_fma(amp2, f2_pack(ex2(ta), ex2(tb)), acc[i]);
    }
  }
}

template <typename OutT>
__device__ __forceinline__ void store4(OutT* p, bool vec, int valid, float4 v) {
  if (vec) {
    if constexpr (sizeof(OutT) == 4) {
      *reinterpret_cast<float4*>(p) = v;
    } else {
      double2* q = reinterpret_cast<double2*>(p);
      q[0] = make_double2((double)v.x, (double)v.y);
      q[1] = make_double2((double)v.z, (double)v.w);
    }
  } else {
    const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < valid) p[j] = (OutT)e[j];
  }
}

// one thread: wait until every tile of the image has arrived, return the image maximum
__device__ __noinline__ float wait_image_peak(const unsigned* arrived, const unsigned* peak, unsigned tiles) {
  const volatile unsigned* arr = arrived;
  const long long t0 = clock64();
  while (*arr < tiles) {
    __nanosleep(64);
    // (cannot happen by the ticket rule; if it ever does, fail loudly instead of hanging the device)
    if (clock64() - t0 > 40000000000ll) __trap();
  }
  __threadfence();
  return __uint_as_float(*(const volatile unsigned*)peak);
}

template <typename OutT>
__global__ void __launch_bounds__(kThreads, 1) tile_kernel(const TileArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  const int N = a.N, total_tiles = a.n * a.tiles;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool vec = (N & 3) == 0 && (reinterpret_cast<uintptr_t>(a.out_x) & 15) == 0 &&
                   (reinterpret_cast<uintptr_t>(a.out_y) & 15) == 0;
  const size_t npx = (size_t)N * N;

  // scale and write a finished tile once its image's maximum is final
  auto flush = [&](int ticket, int pb) {
    const int img = ticket / a.tiles, tile = a.order[ticket];
    const int ty = tile / a.tiles_x, tx = tile - ty * a.tiles_x;
    const int R0 = ty * kTileRows, C0 = tx * kTileCols;
    float inv = 1.f;
    if (!a.direct) {
      if (tid == 0) {
        sm.inv = 1.f / wait_image_peak(a.arrived + img, a.peak + img, (unsigned)a.tiles);
      }
      __syncthreads();
      inv = sm.inv;
    }
    OutT* x = reinterpret_cast<OutT*>(a.out_x) + (size_t)img * npx;
    OutT* y = reinterpret_cast<OutT*>(a.out_y) + (size_t)img * 4 * npx;
    const float* p0 = sm.plane[pb][0];
    for (int e = tid; e < kTilePx / 4; e += kThreads) {
      const int rl = e >> 5, c4 = (e & 31) << 2;
      const int r = R0 + rl, c = C0 + c4;
      if (r >= N || c >= N) continue;
      const float4 a0 = *reinterpret_cast<const float4*>(p0 + 4 * e);
      const float4 a1 = *reinterpret_cast<const float4*>(p0 + kTilePx + 4 * e);
      const float4 a2 = *reinterpret_cast<const float4*>(p0 + 2 * kTilePx + 4 * e);
      float4 s;
      s.x = ((a0.x + a1.x) + a2.x) * inv; s.y = ((a0.y + a1.y) + a2.y) * inv;
      s.z = ((a0.z + a1.z) + a2.z) * inv; s.w = ((a0.w + a1.w) + a2.w) * inv;
      const size_t o = (size_t)r * N + c;
      const int valid = min(4, N - c);
      store4(x + o, vec, valid, s);
      store4(y + o, vec, valid, make_float4(a0.x * inv, a0.y * inv, a0.z * inv, a0.w * inv));
      store4(y + npx + o, vec, valid, make_float4(a1.x * inv, a1.y * inv, a1.z * inv, a1.w * inv));
      store4(y + 2 * npx + o, vec, valid, make_float4(a2.x * inv, a2.y * inv, a2.z * inv, a2.w * inv));
      if (!a.direct)
        store4(y + 3 * npx + o, vec, valid, make_float4(1.f - s.x, 1.f - s.y, 1.f - s.z, 1.f - s.w));
    }
  };

  if (warp == kBuilderWarp) build_next(a, sm.list[0], sm.rlist[0], &sm.info[0], lane, -1);
  __syncthreads();

  int lb = 0, pb = 0;      // list buffer / plane buffer of the current tile
  int prev_ticket = -1;    // finished tile waiting in plane buffer pb ^ 1
  for (;;) {
    ListInfo li = sm.info[lb];
    if (li.ticket >= total_tiles) break;  // (a deferred ticket was replaced before the loop turned)
    const int img = li.ticket / a.tiles, tile = a.order[li.ticket];
    const int ty = tile / a.tiles_x, tx = tile - ty * a.tiles_x;
    const int R0 = ty * kTileRows, C0 = tx * kTileCols;
    const int o2 = a.offs[img * 3 + 1], o3 = a.offs[img * 3 + 2], j1 = a.jet_offs[img + 1];
    float* planes = sm.plane[pb][0];

    for (int round = 0;; ++round) {
      // ---- gather the staged records -----------------------------------------------------
      for (int k = tid; k < li.nsrc; k += kThreads) {
        const float4* src = reinterpret_cast<const float4*>(a.rec + sm.list[lb][k]);
        float4* dst = reinterpret_cast<float4*>(&sm.rec[k]);
        const float4 v0 = __ldg(src), v1 = __ldg(src + 1), v2 = __ldg(src + 2);
        dst[0] = v0; dst[1] = v1; dst[2] = v2;
      }
      for (int k = tid; k < li.nrect; k += kThreads) sm.rect[k] = a.jets[sm.rlist[lb][k]];
      if (tid == 0) sm.task = 0;
      __syncthreads();
#ifdef RSB_DBG
      if (tid == 0) {
        printf("cta %d ticket %d round %d nsrc %d nwide %d nrect %d resume %d %d list:", blockIdx.x, li.ticket, round,
               li.nsrc, li.nwide, li.nrect, li.src_resume, li.rect_resume);
        for (int k = 0; k < li.nsrc; ++k) printf(" %u(A=%g,k2=%g)", sm.list[lb][k], sm.rec[k].A, sm.rec[k].k2);
        printf("\n");
      }
#endif
      // the builder warp prepares the next tile while the others render this one
      const bool last_round = li.src_resume < 0 && li.rect_resume < 0;
      if (warp == kBuilderWarp && last_round)
        build_next(a, sm.list[lb ^ 1], sm.rlist[lb ^ 1], &sm.info[lb ^ 1], lane,
                   (prev_ticket >= 0 && !a.direct) ? prev_ticket / a.tiles : -1);

      // ---- tasks ---------------------------------------------------------------------------
      // 48 tasks: class (wide, narrow, sub-jets) x sub-tile, handed out by a counter.  Queue
      // order: wide (exp-bound) and sub-jet (gather-latency-bound) tasks alternate, so both kinds
      // are in flight on the SM at any time; the cheap narrow tasks come last and even out the end.
      auto next_task = [&]() -> int {
        int t = 0;
        if (lane == 0) t = atomicAdd(&sm.task, 1);
        return __shfl_sync(0xffffffffu, t, 0);
      };
      for (int q = next_task(); q < 3 * kSub; q = next_task()) {
        // q -> task: [0, 32): even = wide, odd = sub-jets of sub-tile q/2; [32, 48): narrow
        const int t = q < 2 * kSub ? ((q & 1) ? 2 * kSub + (q >> 1) : (q >> 1)) : q - kSub;
        const bool is_src = t < kSrcTasks;
        const int cls = t >= kSub ? 1 : 0;                // 0: wide, 1: narrow
        const int u = t & (kSub - 1);
        const int g = u >> 1, half = u & 1;
        const int lo = is_src ? (cls ? li.nwide : 0) : 0;
        const int hi = is_src ? (cls ? li.nsrc : li.nwide) : li.nrect;
        if (round > 0 && lo >= hi) continue;              // nothing to add in a later round
        const int r0 = R0 + g * kSubRows, c0t = C0 + half * kSubCols, c0 = c0t + 2 * lane;
        float* dst = planes + (is_src ? (1 + cls) : 0) * kTilePx + (g * kSubRows) * kTileCols +
                     half * kSubCols + 2 * lane;
        f2 acc[kSubRows];
#pragma unroll
        for (int i = 0; i < kSubRows; ++i) acc[i] = 0ull;
        if (r0 < N && c0t < N) {
          if (is_src) {
            const float r0f = (float)r0, c0f = (float)c0;
            // rectangle of the task with 1.5 pixels of slack
            const float ra = (float)r0 - 1.5f, rb = (float)(r0 + kSubRows - 1) + 1.5f;
            const float ca = (float)c0t - 1.5f, cb = (float)(c0t + kSubCols - 1) + 1.5f;
            for (int k0 = lo; k0 < hi; k0 += 32) {
              bool hit = false;
              if (k0 + lane < hi) {
                const Rec* rk = &sm.rec[k0 + lane];
                const float4 q = *reinterpret_cast<const float4*>(&rk->A);
                const float4 ct = *reinterpret_cast<const float4*>(&rk->irf);
                const float4 kk = *reinterpret_cast<const float4*>(&rk->k2);  // k2 T2s g1 g2
                const float yc = ct.x + ct.y, xc = ct.z + ct.w;
                const float dra = ra - yc, drb = rb - yc, dca = ca - xc, dcb = cb - xc;
                // only the edges that face the centre can hold the maximum of q
                const bool use_c = dca > 0.f || dcb < 0.f, use_r = dra > 0.f || drb < 0.f;
                const float e = dca > 0.f ? dca : dcb, f = dra > 0.f ? dra : drb;
                const float tr = fminf(fmaxf(-kk.z * e, dra), drb);   // column edge dc = e, dr free
                const float qc = (q.z * tr + q.y * e) * tr + q.x * e * e;
                const float tc = fminf(fmaxf(-kk.w * f, dca), dcb);   // row edge dr = f, dc free
                const float qr = (q.x * tc + q.y * f) * tc + q.z * f * f;
                float qmax = fmaxf(use_c ? qc : -3.0e38f, use_r ? qr : -3.0e38f);
                if (!use_c && !use_r) qmax = 0.f;  // the centre lies inside the rectangle
                hit = !(qmax < kk.y);              // (NaN from a degenerate source counts as touching)
              }
              unsigned hits = __ballot_sync(0xffffffffu, hit);
              if (!hits) continue;
              // Coefficients of the hit about to be rendered; the next hit's are requested before
              // this one's arithmetic (the warp barrier keeps the loads above it).  Unrolled by
              // two so that the two coefficient sets swap roles without register moves.
              auto fetch = [&](int b, float4& q_, float4& c_, float2& k_) {
                const Rec* rn = &sm.rec[k0 + max(b, 0)];
                q_ = *reinterpret_cast<const float4*>(&rn->A);
                c_ = *reinterpret_cast<const float4*>(&rn->irf);
                k_ = *reinterpret_cast<const float2*>(&rn->k2);
              };
              float4 q, ct, qb, cb2;
              float2 kk, kb;
              fetch(lowest_bit(hits), q, ct, kk);
              for (;;) {
                hits &= hits - 1;
                fetch(lowest_bit(hits), qb, cb2, kb);
                __syncwarp();
                source_math(q, ct, kk, r0f, c0f, acc);
                if (!hits) break;
                hits &= hits - 1;
                fetch(lowest_bit(hits), q, ct, kk);
                __syncwarp();
                source_math(qb, cb2, kb, r0f, c0f, acc);
                if (!hits) break;
              }
            }
          } else {
            // class 0: paste the sub-jets that overlap the task's rectangle (survey.py:61-76).
            // Branch-free gathers from clamped addresses (16 independent loads per hit, masked
            // afterwards), and the next hit's loads are issued before this hit's additions.
            float a0[kSubRows], a1[kSubRows];
#pragma unroll
            for (int i = 0; i < kSubRows; ++i) a0[i] = a1[i] = 0.f;
            for (int k0 = 0; k0 < hi; k0 += 32) {
              bool hit = false;
              if (k0 + lane < hi) {
                const JetRect jr = sm.rect[k0 + lane];
                hit = !(jr.row0 > r0 + kSubRows - 1 || jr.row0 + jr.size <= r0 ||
                        jr.col0 > c0t + kSubCols - 1 || jr.col0 + jr.size <= c0t);
              }
              unsigned hits = __ballot_sync(0xffffffffu, hit);
              if (!hits) continue;
              float v0[kSubRows], v1[kSubRows];   // pixels of the hit in flight
              int fr = 0, fc = 0, fs = 0;         // ... its first row / column relative to the task, its size
              auto gather = [&](int b) {
                const JetRect h = sm.rect[k0 + max(b, 0)];
                const float* jp = a.jet_pix + h.pix;
                fr = r0 - h.row0; fc = c0 - h.col0; fs = h.size;
                const int jc0 = min(max(fc, 0), fs - 1), jc1 = min(max(fc + 1, 0), fs - 1);
#pragma unroll
                for (int i = 0; i < kSubRows; ++i) {
                  const float* rowp = jp + (size_t)min(max(fr + i, 0), fs - 1) * h.pitch;
                  v0[i] = __ldg(rowp + jc0);
                  v1[i] = __ldg(rowp + jc1);
                }
              };
              gather(lowest_bit(hits));
              while (hits) {
                hits &= hits - 1;
                float w0[kSubRows], w1[kSubRows];
                const int gr = fr, gc = fc, gs = fs;
#pragma unroll
                for (int i = 0; i < kSubRows; ++i) { w0[i] = v0[i]; w1[i] = v1[i]; }
                if (hits) gather(lowest_bit(hits));   // (warp-uniform)
                const bool in0 = gc >= 0 && gc < gs, in1 = gc + 1 >= 0 && gc + 1 < gs;
#pragma unroll
                for (int i = 0; i < kSubRows; ++i) {
                  const bool rin = gr + i >= 0 && gr + i < gs;
                  a0[i] += (rin && in0) ? w0[i] : 0.f;
                  a1[i] += (rin && in1) ? w1[i] : 0.f;
                }
              }
            }
#pragma unroll
            for (int i = 0; i < kSubRows; ++i) acc[i] = f2_pack(a0[i], a1[i]);
          }
        }
        if (round == 0) {
#pragma unroll
          for (int i = 0; i < kSubRows; ++i) *reinterpret_cast<f2*>(dst + i * kTileCols) = acc[i];
        } else {
#pragma unroll
          for (int i = 0; i < kSubRows; ++i) {
            f2* d2 = reinterpret_cast<f2*>(dst + i * kTileCols);
            *d2 = f2_add(*d2, acc[i]);
          }
        }
      }
      __syncthreads();  // every task of this round is in the planes; rec / rect / list are free
      if (last_round) break;
      if (warp == 0)
        build_round(a, li.ticket, li.src_resume < 0 ? o3 : li.src_resume, o2, o3,
                    li.rect_resume < 0 ? j1 : li.rect_resume, j1, R0, C0, sm.list[lb], sm.rlist[lb],
                    &sm.info[lb], lane);
      __syncthreads();
      li = sm.info[lb];
    }

    // ---- tile maximum of the plane sum -> per-image maximum; the tile arrives -------------
    {
      float m = 0.f;
      for (int e = tid; e < kTilePx / 4; e += kThreads) {
        const int rl = e >> 5, c4 = (e & 31) << 2;
        const int r = R0 + rl, c = C0 + c4;
        if (r >= N || c >= N) continue;
        const float4 a0 = *reinterpret_cast<const float4*>(planes + 4 * e);
        const float4 a1 = *reinterpret_cast<const float4*>(planes + kTilePx + 4 * e);
        const float4 a2 = *reinterpret_cast<const float4*>(planes + 2 * kTilePx + 4 * e);
        m = fmaxf(m, (a0.x + a1.x) + a2.x);
        if (c + 1 < N) m = fmaxf(m, (a0.y + a1.y) + a2.y);
        if (c + 2 < N) m = fmaxf(m, (a0.z + a1.z) + a2.z);
        if (c + 3 < N) m = fmaxf(m, (a0.w + a1.w) + a2.w);
      }
      m = warp_max(m);
      if (lane == 0) sm.red[warp] = m;
      __syncthreads();
      if (tid == 0) {
        float mm = 0.f;
        for (int w = 0; w < kWarps; ++w) mm = fmaxf(mm, sm.red[w]);
        // values >= 0: float order = unsigned order
        atomicMax(a.peak + img, __float_as_uint(fmaxf(mm, 0.f)));
        __threadfence();
        atomicAdd(a.arrived + img, 1u);
      }
    }
    if (a.direct) {
      __syncthreads();
      flush(li.ticket, pb);
      __syncthreads();  // the planes are free again; the next tile's list is visible
      lb ^= 1;
      continue;
    }
    // ---- flush the tile finished one step ago (its image has had a whole tile time) --------
    if (prev_ticket >= 0) flush(prev_ticket, pb ^ 1);
    prev_ticket = li.ticket;
    pb ^= 1;
    lb ^= 1;
    __syncthreads();  // the next tile's list (builder warp) is visible; buffers swapped
    if (sm.info[lb].ticket == kDeferred) {  // (uniform) the ticket the builder had to leave alone
      __syncthreads();
      if (warp == kBuilderWarp) build_next(a, sm.list[lb], sm.rlist[lb], &sm.info[lb], lane, -1);
      __syncthreads();
    }
  }
  if (prev_ticket >= 0 && !a.direct) flush(prev_ticket, pb ^ 1);
}

// ---- tile order: most expensive tiles of an image first ------------------------------------
// An image is complete (its maximum final, its tiles free to leave the chip) when its LAST
// tile is done; handing out the cheap border tiles last keeps that moment close to the last
// ticket.  Cost proxy: pixels of the tile inside the sources' boxes (+ a quarter for pasted
// sub-jet pixels).  One warp per tile.
__global__ void __launch_bounds__(128) cost_kernel(const TileArgs a, unsigned* __restrict__ cost) {
  const int lane = threadIdx.x & 31, tile = blockIdx.x * 4 + (threadIdx.x >> 5), img = blockIdx.y;
  if (tile >= a.tiles) return;
  const int ty = tile / a.tiles_x, tx = tile - ty * a.tiles_x;
  const int R0 = ty * kTileRows, C0 = tx * kTileCols, R1 = R0 + kTileRows - 1, C1 = C0 + kTileCols - 1;
  const int o1 = a.offs[img * 3], o3 = a.offs[img * 3 + 2];
  unsigned sum = 0u;
  for (int s = o1 + lane; s < o3; s += 32) {
    const int4 bb = a.bbox[s];
    const int h = min(bb.y, R1) - max(bb.x, R0) + 1, w = min(bb.w, C1) - max(bb.z, C0) + 1;
    if (h > 0 && w > 0) sum += (unsigned)(h * w);
  }
  unsigned rsum = 0u;
  for (int j = a.jet_offs[img] + lane; j < a.jet_offs[img + 1]; j += 32) {
    const JetRect jr = a.jets[j];
    const int h = min(jr.row0 + jr.size - 1, R1) - max(jr.row0, R0) + 1;
    const int w = min(jr.col0 + jr.size - 1, C1) - max(jr.col0, C0) + 1;
    if (h > 0 && w > 0) rsum += (unsigned)(h * w);
  }
  sum += rsum >> 2;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) cost[(size_t)img * a.tiles + tile] = sum;
}

// one CTA per image: order[rank] = tile, rank by descending cost (ties: tile index)
__global__ void __launch_bounds__(256) order_kernel(const unsigned* __restrict__ cost, int tiles,
                                                    int* __restrict__ order) {
  extern __shared__ unsigned c_sm[];
  const int img = blockIdx.x;
  for (int j = threadIdx.x; j < tiles; j += blockDim.x) c_sm[j] = cost[(size_t)img * tiles + j];
  __syncthreads();
  for (int j = threadIdx.x; j < tiles; j += blockDim.x) {
    const unsigned cj = c_sm[j];
    int rank = 0;
    for (int k = 0; k < tiles; ++k) rank += (c_sm[k] > cj || (c_sm[k] == cj && k < j)) ? 1 : 0;
    order[(size_t)img * tiles + rank] = j;
  }
}

__global__ void identity_order_kernel(int tiles, size_t total, int* __restrict__ order) {
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (size_t)gridDim.x * blockDim.x)
    order[e] = (int)(e % (size_t)tiles);
}

// direct mode only: x *= 1/max, y[0..2] *= 1/max, y[3] = 1 - x, in place
template <typename OutT>
__global__ void __launch_bounds__(256) norm_inplace_kernel(const unsigned* __restrict__ peak, int N,
                                                           OutT* __restrict__ out_x, OutT* __restrict__ out_y) {
  const size_t npx = (size_t)N * N;
  const int img = blockIdx.y;
  const float inv = 1.f / __uint_as_float(peak[img]);
  OutT* x = out_x + (size_t)img * npx;
  OutT* y = out_y + (size_t)img * 4 * npx;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < npx; e += (size_t)gridDim.x * blockDim.x) {
    const float a0 = (float)y[e], a1 = (float)y[npx + e], a2 = (float)y[2 * npx + e];
    const float s = ((a0 + a1) + a2) * inv;
    x[e] = (OutT)s;
    y[e] = (OutT)(a0 * inv);
    y[npx + e] = (OutT)(a1 * inv);
    y[2 * npx + e] = (OutT)(a2 * inv);
    y[3 * npx + e] = (OutT)(1.f - s);
  }
}

}  // namespace splat
}  // namespace rsb

using namespace rsb;
using namespace rsb::splat;

static size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

// The CTAs of tile_kernel wait for each other (a tile leaves the chip when its image is complete), which is
// safe while the kernel owns enough SMs; two such kernels running side by side on different streams could each
// hold too few.  So tile kernels of one device are chained: each launch waits for the previous one's event.
#include <mutex>
namespace {
std::mutex g_chain_mu;
cudaEvent_t g_chain_ev[64] = {};
}  // namespace

extern "C" {

int rsb_survey_workspace_bytes(int n, int size, int total_sources, size_t* bytes) {
  RSB_ARG(bytes != nullptr && n >= 0 && size >= 2 && total_sources >= 0,
          "rsb_survey_workspace_bytes: bad arguments");
  const size_t s = (size_t)(total_sources > 0 ? total_sources : 1);
  const size_t tiles = (size_t)((size + kTileCols - 1) / kTileCols) * ((size + kTileRows - 1) / kTileRows);
  *bytes = align256((2 * (size_t)n + 1) * sizeof(unsigned)) + align256(s * sizeof(Rec)) +
           align256(s * sizeof(int4)) + 2 * align256((size_t)n * tiles * sizeof(unsigned));
  return 0;
}

int rsb_survey_render(const double* src_params, const int32_t* src_offs, int total_sources,
                      const void* jet_rects, const int32_t* jet_offs, const float* jet_pix, int n,
                      int size, unsigned flags, void* out_x, void* out_y, void* workspace,
                      size_t workspace_bytes, void* stream) {
  RSB_ARG(n >= 0 && size >= 2 && size <= 16384, "rsb_survey_render: bad n/size (%d, %d)", n, size);
  if (n == 0) return 0;
  RSB_ARG(src_offs && jet_offs && out_x && out_y && workspace, "rsb_survey_render: NULL buffer");
  RSB_ARG(total_sources == 0 || src_params, "rsb_survey_render: NULL source parameters");
  size_t need = 0;
  if (int e = rsb_survey_workspace_bytes(n, size, total_sources, &need)) return e;
  RSB_ARG(workspace_bytes >= need, "rsb_survey_render: workspace too small (%zu < %zu bytes)",
          workspace_bytes, need);
  const int tiles_x = (size + kTileCols - 1) / kTileCols, tiles_y = (size + kTileRows - 1) / kTileRows;
  RSB_ARG((long long)n * tiles_x * tiles_y < (1ll << 30), "rsb_survey_render: too many tiles in one call");
  cudaStream_t st = (cudaStream_t)stream;
  char* w = (char*)workspace;
  unsigned* ctl = (unsigned*)w;  // peak[n], arrived[n], ticket
  const size_t ctl_bytes = (2 * (size_t)n + 1) * sizeof(unsigned);
  size_t off = align256(ctl_bytes);
  Rec* rec = (Rec*)(w + off);
  off += align256((size_t)(total_sources > 0 ? total_sources : 1) * sizeof(Rec));
  int4* bbox = (int4*)(w + off);
  off += align256((size_t)(total_sources > 0 ? total_sources : 1) * sizeof(int4));
  const size_t ntile = (size_t)n * tiles_x * tiles_y;
  unsigned* cost = (unsigned*)(w + off);
  off += align256(ntile * sizeof(unsigned));
  int* order = (int*)(w + off);
  RSB_CUDA(cudaMemsetAsync(ctl, 0, ctl_bytes, st));
  if (total_sources > 0) {
    // tau: 1e-9 of the largest amplitude any source can have (amp ~ U(0,1), sub-jets peak at 1)
    prep_kernel<<<(total_sources + 255) / 256, 256, 0, st>>>(src_params, total_sources, size, 1e-9, rec, bbox);
    RSB_LAUNCHED();
  }
  TileArgs a;
  a.rec = rec; a.bbox = bbox; a.offs = src_offs; a.jets = (const JetRect*)jet_rects; a.jet_offs = jet_offs;
  a.jet_pix = jet_pix; a.peak = ctl; a.arrived = ctl + n; a.ticket = ctl + 2 * (size_t)n;
  a.out_x = out_x; a.out_y = out_y; a.n = n; a.N = size; a.tiles_x = tiles_x; a.tiles = tiles_x * tiles_y;
  const size_t smem = sizeof(Smem);
  int grid = sm_count();
  if ((long long)grid > (long long)n * a.tiles) grid = n * a.tiles;
  a.direct = a.tiles > grid ? 1 : 0;
  a.order = order;
  int dev = 0;
  RSB_CUDA(cudaGetDevice(&dev));
  RSB_ARG(dev >= 0 && dev < 64, "rsb_survey_render: device index %d out of range", dev);
  std::lock_guard<std::mutex> chain(g_chain_mu);   // (held until this call's kernel and event are enqueued)
  if (!g_chain_ev[dev]) RSB_CUDA(cudaEventCreateWithFlags(&g_chain_ev[dev], cudaEventDisableTiming));
  else RSB_CUDA(cudaStreamWaitEvent(st, g_chain_ev[dev], 0));
  if (a.direct) {
    identity_order_kernel<<<64, 256, 0, st>>>(a.tiles, ntile, order);
    RSB_LAUNCHED();
  } else {
    cost_kernel<<<dim3((a.tiles + 3) / 4, n), 128, 0, st>>>(a, cost);
    RSB_LAUNCHED();
    order_kernel<<<n, 256, (size_t)a.tiles * sizeof(unsigned), st>>>(cost, a.tiles, order);
    RSB_LAUNCHED();
  }
  if (flags & RSB_JET_OUT_F64) {
    RSB_CUDA(cudaFuncSetAttribute(tile_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tile_kernel<double><<<grid, kThreads, smem, st>>>(a);
  } else {
    RSB_CUDA(cudaFuncSetAttribute(tile_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    tile_kernel<float><<<grid, kThreads, smem, st>>>(a);
  }
  RSB_LAUNCHED();
  RSB_CUDA(cudaEventRecord(g_chain_ev[dev], st));
  if (a.direct) {
    int bx = (int)(((size_t)size * size + 255) / 256);
    if (bx > 2048) bx = 2048;
    for (int i0 = 0; i0 < n; i0 += 65535) {
      const int cnt = (n - i0) < 65535 ? (n - i0) : 65535;
      const size_t px = (size_t)size * size;
      if (flags & RSB_JET_OUT_F64)
        norm_inplace_kernel<double><<<dim3(bx, cnt), 256, 0, st>>>(a.peak + i0, size, (double*)out_x + i0 * px,
                                                                   (double*)out_y + (size_t)i0 * 4 * px);
      else
        norm_inplace_kernel<float><<<dim3(bx, cnt), 256, 0, st>>>(a.peak + i0, size, (float*)out_x + i0 * px,
                                                                  (float*)out_y + (size_t)i0 * 4 * px);
      RSB_LAUNCHED();
    }
  }
  return 0;
}

}  // extern "C"
