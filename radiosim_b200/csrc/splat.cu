// Tiled Gaussian splat for survey images (many components on large grids).
//
// Reference: src/radiosim/survey/survey.py:49-122 -- every source is evaluated on the full
// N x N grid (gauss.py:50-97) and added to its class plane; the image is the plane sum
// divided by its maximum, the last plane is 1 - image.
//
// Here:
//   prep_kernel   : float64 source parameters -> float32 conic coefficient records with the
//                   bounding box of the > tau footprint (one thread per source).
//   tile_kernel   : one CTA per 128 x 128 image tile.  The image's coefficient records stream
//                   through shared memory in chunks of 512 (1-D bulk async copies = TMA
//                   engine, double buffered on mbarriers).  Each warp owns four 8 x 64
//                   sub-tiles (16 pixels per lane each, accumulators in registers) and culls
//                   the staged chunk against them with one ballot per 32 sources; a hit costs
//                   one MUFU.EX2 + 1.5 FFMA2-class instructions per pixel.  Sub-jets (class
//                   0), rendered beforehand by the jet kernel, are pasted with the reference's
//                   clipping.  Unnormalised planes go to a scratch buffer once, the tile
//                   maximum of the plane sum to a per-image atomicMax.
//   norm_kernel   : planes * 1/max, image, background plane -- single pass, 128-bit accesses.
#include "jet_common.cuh"

namespace rsb {
namespace splat {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kTile = 128;        // CTA tile edge
constexpr int kSubRows = 8;       // warp sub-tile: 8 rows x 64 columns
constexpr int kSubCols = 64;
constexpr int kSubPerWarp = (kTile / kSubRows) * (kTile / kSubCols) / kWarps;  // 4
constexpr int kChunk = 512;       // coefficient records per staged chunk (24 KB)

using jet::Coef;

struct JetRect {  // one pasted sub-jet (survey.py:56-76)
  long long pix;  // offset of its first pixel in the packed jet buffer
  int size, row0, col0, pad;
};

// ---- prep: parameters -> coefficient records -------------------------------------------
__global__ void prep_kernel(const double* __restrict__ params, int total, int N, double tau,
                            Coef* __restrict__ out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= total) return;
  const double* p = params + (size_t)k * 6;
  const double amp = p[0], x = p[1], y = p[2], sx = p[3], sy = p[4];
  Coef c;
  c.A = c.B = c.C = 0.f; c.amp = 0.f; c.irf = c.fr = c.icf = c.fc = 0.f;
  c.r0 = 1; c.r1 = 0; c.c0 = 1; c.c1 = 0;  // empty box: never hit
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
    c.amp = (float)amp;
    const double yr = rint(y), xr = rint(x);
    c.irf = (float)yr; c.fr = (float)(y - yr);
    c.icf = (float)xr; c.fc = (float)(x - xr);
    const double T2 = 2.0 * log(fabs(amp) / tau), det = Mrr * Mcc - Mrc * Mrc;
    const double er = sqrt(T2 * Mcc / det), ec = sqrt(T2 * Mrr / det);
    if (isfinite(er) && isfinite(ec)) {
      const double lo_r = fmax(floor(y - er) - 1.0, 0.0), hi_r = fmin(ceil(y + er) + 1.0, N - 1.0);
      const double lo_c = fmax(floor(x - ec) - 1.0, 0.0), hi_c = fmin(ceil(x + ec) + 1.0, N - 1.0);
      if (lo_r <= hi_r && lo_c <= hi_c) {
        c.r0 = (int)lo_r; c.r1 = (int)hi_r; c.c0 = (int)lo_c; c.c1 = (int)hi_c;
      }
    } else {  // degenerate widths: never cull
      c.r0 = 0; c.r1 = N - 1; c.c0 = 0; c.c1 = N - 1;
    }
  }
  out[k] = c;
}

struct TileArgs {
  const Coef* coef;          // [total] records, grouped per image: class 1 then class 2
  const int32_t* offs;       // [n][3]: first class-1 record, first class-2 record, end
  const JetRect* jets;       // [total jets]
  const int32_t* jet_offs;   // [n + 1]
  const float* jet_pix;      // packed sub-jet images (already normalised)
  float* planes;             // [n][3][N][N] unnormalised class planes (scratch)
  unsigned* peak;            // [n] float bits of max(sum of planes), zero-initialised
  int n, N;
};

struct __align__(16) TileSmem {
  Coef chunk[2][kChunk];
  uint64_t bar[2];
};

__global__ void __launch_bounds__(kThreads) tile_kernel(const TileArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  TileSmem& sm = *reinterpret_cast<TileSmem*>(smem_raw);
  const int N = a.N, img = blockIdx.y;
  const int tiles_x = (N + kTile - 1) / kTile;
  const int ty = blockIdx.x / tiles_x, tx = blockIdx.x - ty * tiles_x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int R0 = ty * kTile, C0 = tx * kTile;

  if (tid == 0) {
    mbar_init(&sm.bar[0], 1);
    mbar_init(&sm.bar[1], 1);
    fence_mbar_init();
  }
  __syncthreads();

  // this warp's sub-tiles: s -> (row group, column half)
  int sr0[kSubPerWarp], sc0[kSubPerWarp];
#pragma unroll
  for (int s = 0; s < kSubPerWarp; ++s) {
    const int id = warp * kSubPerWarp + s;            // 0..31
    sr0[s] = R0 + (id >> 1) * kSubRows;
    sc0[s] = C0 + (id & 1) * kSubCols;
  }
  f2 total[kSubPerWarp][kSubRows];
#pragma unroll
  for (int s = 0; s < kSubPerWarp; ++s)
#pragma unroll
    for (int i = 0; i < kSubRows; ++i) total[s][i] = 0ull;

  const int o1 = a.offs[img * 3], o2 = a.offs[img * 3 + 1], o3 = a.offs[img * 3 + 2];
  uint32_t phase[2] = {0u, 0u};
  float* plane_base = a.planes + (size_t)img * 3 * N * N;

  for (int cls = 1; cls <= 2; ++cls) {
    const int first = cls == 1 ? o1 : o2, last = cls == 1 ? o2 : o3;
    f2 acc[kSubPerWarp][kSubRows];
#pragma unroll
    for (int s = 0; s < kSubPerWarp; ++s)
#pragma unroll
      for (int i = 0; i < kSubRows; ++i) acc[s][i] = 0ull;
    const int nchunks = (last - first + kChunk - 1) / kChunk;
    if (tid == 0 && nchunks > 0) {
      const int cnt = min(kChunk, last - first);
      mbar_expect_tx(&sm.bar[0], (uint32_t)cnt * (uint32_t)sizeof(Coef));
      bulk_g2s(sm.chunk[0], a.coef + first, (uint32_t)cnt * (uint32_t)sizeof(Coef), &sm.bar[0]);
    }
    for (int ch = 0; ch < nchunks; ++ch) {
      const int b = ch & 1;
      const int cnt = min(kChunk, last - first - ch * kChunk);
      if (tid == 0 && ch + 1 < nchunks) {  // stage the next chunk while this one is consumed
        const int ncnt = min(kChunk, last - first - (ch + 1) * kChunk);
        fence_proxy_async();
        mbar_expect_tx(&sm.bar[b ^ 1], (uint32_t)ncnt * (uint32_t)sizeof(Coef));
        bulk_g2s(sm.chunk[b ^ 1], a.coef + first + (size_t)(ch + 1) * kChunk,
                 (uint32_t)ncnt * (uint32_t)sizeof(Coef), &sm.bar[b ^ 1]);
      }
      mbar_wait(&sm.bar[b], phase[b]);
      phase[b] ^= 1u;
      const Coef* coef = sm.chunk[b];
#pragma unroll
      for (int s = 0; s < kSubPerWarp; ++s) {
        const int r0 = sr0[s], c0t = sc0[s];
        if (r0 >= N || c0t >= N) continue;
        const float r0f = (float)r0, c0f = (float)(c0t + 2 * lane);
        f2 part[kSubRows];  // this chunk's contribution: keeps float32 sums short
#pragma unroll
        for (int i = 0; i < kSubRows; ++i) part[i] = 0ull;
        for (int k0 = 0; k0 < cnt; k0 += 32) {
          bool hit = false;
          if (k0 + lane < cnt) {
            const int4 bb = *reinterpret_cast<const int4*>(&coef[k0 + lane].r0);
            hit = !(r0 > bb.y || r0 + kSubRows - 1 < bb.x || c0t > bb.w || c0t + kSubCols - 1 < bb.z);
          }
          unsigned hits = __ballot_sync(0xffffffffu, hit);
          while (hits) {
            const int k = k0 + __ffs(hits) - 1;
            hits &= hits - 1;
            const float4 q = *reinterpret_cast<const float4*>(&coef[k].A);
            const float4 ct = *reinterpret_cast<const float4*>(&coef[k].irf);
            const float d0 = (c0f - ct.z) - ct.w;
            const f2 d = f2_pack(d0, d0 + 1.f);
            const f2 Q = f2_mul(f2_pack(q.y, q.y), d);
            const f2 Pq = f2_mul(f2_mul(f2_pack(q.x, q.x), d), d);
            const f2 amp2 = f2_pack(q.w, q.w);
            const float dr0 = (r0f - ct.x) - ct.y;
#pragma unroll
            for (int i = 0; i < kSubRows; i += 2) {
              const f2 drp = f2_pack(dr0 + (float)i, dr0 + (float)(i + 1));
              const f2 Cqp = f2_mul(f2_mul(f2_pack(q.z, q.z), drp), drp);
              float dra, drb, Cqa, Cqb, ta, tb;
              f2_unpack(drp, dra, drb);
              f2_unpack(Cqp, Cqa, Cqb);
              f2_unpack(f2_fma(f2_pack(dra, dra), Q, f2_add(Pq, f2_pack(Cqa, Cqa))), ta, tb);
              part[i] = f2_fma(amp2, f2_pack(ex2(ta), ex2(tb)), part[i]);
              f2_unpack(f2_fma(f2_pack(drb, drb), Q, f2_add(Pq, f2_pack(Cqb, Cqb))), ta, tb);
              part[i + 1] = f2_fma(amp2, f2_pack(ex2(ta), ex2(tb)), part[i + 1]);
            }
          }
        }
#pragma unroll
        for (int i = 0; i < kSubRows; ++i) acc[s][i] = f2_add(acc[s][i], part[i]);
      }
      __syncthreads();  // everyone is done with chunk b before it is refilled
    }
    // write this class plane (unnormalised) and add it to the running sum
    float* plane = plane_base + (size_t)cls * N * N;
#pragma unroll
    for (int s = 0; s < kSubPerWarp; ++s) {
      const int c = sc0[s] + 2 * lane;
#pragma unroll
      for (int i = 0; i < kSubRows; ++i) {
        const int r = sr0[s] + i;
        total[s][i] = f2_add(total[s][i], acc[s][i]);
        if (r < N && c < N) {
          float v0, v1;
          f2_unpack(acc[s][i], v0, v1);
          if (c + 1 < N && ((N & 1) == 0)) *reinterpret_cast<float2*>(plane + (size_t)r * N + c) = make_float2(v0, v1);
          else {
            plane[(size_t)r * N + c] = v0;
            if (c + 1 < N) plane[(size_t)r * N + c + 1] = v1;
          }
        }
      }
    }
  }

  // class 0: paste the sub-jets that overlap each sub-tile (survey.py:61-76)
  {
    const int j0 = a.jet_offs[img], j1 = a.jet_offs[img + 1];
    float* plane = plane_base;
#pragma unroll
    for (int s = 0; s < kSubPerWarp; ++s) {
      const int r0 = sr0[s], c0t = sc0[s], c = c0t + 2 * lane;
      float acc0[kSubRows], acc1[kSubRows];
#pragma unroll
      for (int i = 0; i < kSubRows; ++i) acc0[i] = acc1[i] = 0.f;
      if (r0 < N && c0t < N) {
        for (int k0 = j0; k0 < j1; k0 += 32) {
          bool hit = false;
          JetRect jr;
          if (k0 + lane < j1) {
            jr = a.jets[k0 + lane];
            hit = !(jr.row0 > r0 + kSubRows - 1 || jr.row0 + jr.size <= r0 ||
                    jr.col0 > c0t + kSubCols - 1 || jr.col0 + jr.size <= c0t);
          }
          unsigned hits = __ballot_sync(0xffffffffu, hit);
          while (hits) {
            const int src = __ffs(hits) - 1;
            hits &= hits - 1;
            const long long pix = __shfl_sync(0xffffffffu, jr.pix, src);
            const int js = __shfl_sync(0xffffffffu, jr.size, src);
            const int jr0 = __shfl_sync(0xffffffffu, jr.row0, src);
            const int jc0 = __shfl_sync(0xffffffffu, jr.col0, src);
            const float* jp = a.jet_pix + pix;
            const int jc = c - jc0;
#pragma unroll
            for (int i = 0; i < kSubRows; ++i) {
              const int ji = r0 + i - jr0;
              if (ji >= 0 && ji < js) {
                if (jc >= 0 && jc < js) acc0[i] += jp[(size_t)ji * js + jc];
                if (jc + 1 >= 0 && jc + 1 < js) acc1[i] += jp[(size_t)ji * js + jc + 1];
              }
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < kSubRows; ++i) {
        const int r = r0 + i;
        total[s][i] = f2_add(total[s][i], f2_pack(acc0[i], acc1[i]));
        if (r < N && c < N) {
          plane[(size_t)r * N + c] = acc0[i];
          if (c + 1 < N) plane[(size_t)r * N + c + 1] = acc1[i];
        }
      }
    }
  }

  // tile maximum of the plane sum -> per-image maximum (values >= 0: float order = uint order)
  float m = 0.f;
#pragma unroll
  for (int s = 0; s < kSubPerWarp; ++s) {
    const int c = sc0[s] + 2 * lane;
#pragma unroll
    for (int i = 0; i < kSubRows; ++i) {
      float v0, v1;
      f2_unpack(total[s][i], v0, v1);
      if (sr0[s] + i < N && c < N) m = fmaxf(m, v0);
      if (sr0[s] + i < N && c + 1 < N) m = fmaxf(m, v1);
    }
  }
  m = warp_max(m);
  if (lane == 0) atomicMax(a.peak + img, __float_as_uint(fmaxf(m, 0.f)));
}

// ---- normalise: planes /= max, image = sum / max, background = 1 - image ---------------
template <typename OutT>
__global__ void __launch_bounds__(256) norm_kernel(const float* __restrict__ planes,
                                                   const unsigned* __restrict__ peak, int n, int N,
                                                   OutT* __restrict__ out_x, OutT* __restrict__ out_y) {
  const size_t npx = (size_t)N * N;
  const int img = blockIdx.y;
  const float inv = 1.f / __uint_as_float(peak[img]);
  const float* p = planes + (size_t)img * 3 * npx;
  OutT* x = out_x + (size_t)img * npx;
  OutT* y = out_y + (size_t)img * 4 * npx;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < npx;
       e += (size_t)gridDim.x * blockDim.x) {
    const float a0 = p[e], a1 = p[npx + e], a2 = p[2 * npx + e];
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

extern "C" {

int rsb_survey_workspace_bytes(int n, int size, int total_sources, size_t* bytes) {
  RSB_ARG(bytes != nullptr && n >= 0 && size >= 2 && total_sources >= 0,
          "rsb_survey_workspace_bytes: bad arguments");
  size_t b = (size_t)n * 3 * size * size * sizeof(float);
  b = (b + 255) & ~(size_t)255;
  b += ((size_t)n * sizeof(unsigned) + 255) & ~(size_t)255;
  b += (size_t)(total_sources > 0 ? total_sources : 1) * sizeof(Coef);
  *bytes = b;
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
  cudaStream_t st = (cudaStream_t)stream;
  char* w = (char*)workspace;
  float* planes = (float*)w;
  size_t off = ((size_t)n * 3 * size * size * sizeof(float) + 255) & ~(size_t)255;
  unsigned* peak = (unsigned*)(w + off);
  off += ((size_t)n * sizeof(unsigned) + 255) & ~(size_t)255;
  Coef* coef = (Coef*)(w + off);
  RSB_CUDA(cudaMemsetAsync(peak, 0, (size_t)n * sizeof(unsigned), st));
  if (total_sources > 0) {
    // tau: 1e-9 of the largest amplitude any source can have (amp ~ U(0,1), sub-jets peak at 1)
    prep_kernel<<<(total_sources + 255) / 256, 256, 0, st>>>(src_params, total_sources, size, 1e-9, coef);
    RSB_LAUNCHED();
  }
  TileArgs a;
  a.coef = coef; a.offs = src_offs; a.jets = (const JetRect*)jet_rects; a.jet_offs = jet_offs;
  a.jet_pix = jet_pix; a.planes = planes; a.peak = peak; a.n = n; a.N = size;
  const int tiles = ((size + kTile - 1) / kTile) * ((size + kTile - 1) / kTile);
  const size_t smem = sizeof(TileSmem);
  RSB_CUDA(cudaFuncSetAttribute(tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  for (int i0 = 0; i0 < n; i0 += 65535) {
    const int cnt = (n - i0) < 65535 ? (n - i0) : 65535;
    TileArgs b = a;
    b.offs += (size_t)i0 * 3; b.jet_offs += i0; b.planes += (size_t)i0 * 3 * size * size; b.peak += i0;
    b.n = cnt;
    tile_kernel<<<dim3(tiles, cnt), kThreads, smem, st>>>(b);
    RSB_LAUNCHED();
    int bx = (int)(((size_t)size * size + 255) / 256);
    if (bx > 1024) bx = 1024;
    const size_t esz = (flags & RSB_JET_OUT_F64) ? 8 : 4;
    void* ox = (char*)out_x + (size_t)i0 * size * size * esz;
    void* oy = (char*)out_y + (size_t)i0 * 4 * size * size * esz;
    if (flags & RSB_JET_OUT_F64)
      norm_kernel<double><<<dim3(bx, cnt), 256, 0, st>>>(b.planes, b.peak, cnt, size, (double*)ox, (double*)oy);
    else
      norm_kernel<float><<<dim3(bx, cnt), 256, 0, st>>>(b.planes, b.peak, cnt, size, (float*)ox, (float*)oy);
    RSB_LAUNCHED();
  }
  return 0;
}

}  // extern "C"
