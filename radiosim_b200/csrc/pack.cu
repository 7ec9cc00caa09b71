// Native host-side parameter packer for jets: the reference's RNG chain
// (src/radiosim/jets/jet.py:54-125,168) replayed draw for draw on numpy's *legacy* global
// stream, in C++ instead of Python.  Host code only -- no kernel in this file.
//
// The legacy RandomState is MT19937 plus four derived distributions, restated here from
// numpy's published algorithm (numpy/random/src/mt19937/mt19937.c, legacy-distributions.c,
// distributions.c; numpy 2.x):
//   random_sample : (a >> 5, b >> 6) -> (a * 2^26 + b) / 2^53            (two 32-bit outputs)
//   uniform(l, h) : l + (h - l) * random_sample
//   normal(m, s)  : m + s * legacy_gauss; legacy_gauss = polar Box-Muller that returns f*x2 and
//                   caches f*x1 for the next call (has_gauss / gauss of the state tuple)
//   randint(l, h) : l + masked rejection of one 32-bit output against (h - 1 - l)
// The caller passes the state tuple of np.random.get_state() and writes the updated one back
// with np.random.set_state(), so the stream continues exactly where the reference's would.
//
// Elementary functions are libm's; numpy's array loops may use SIMD variants that differ in the
// last bit, so parameters agree with the Python packer to ~1 ulp (tests/test_packer.py), the
// stream position and every branch exactly.
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <new>
#include <thread>
#include <vector>

#include "common.cuh"

namespace rsb {
namespace pack {

struct Legacy {
  uint32_t* key;  // [624]
  int pos;
  int has_gauss;
  double gauss;

  void refill() {
    constexpr int N = 624, M = 397;
    constexpr uint32_t MATRIX_A = 0x9908b0dfu, UPPER = 0x80000000u, LOWER = 0x7fffffffu;
    uint32_t y;
    int i;
    for (i = 0; i < N - M; i++) {
      y = (key[i] & UPPER) | (key[i + 1] & LOWER);
      key[i] = key[i + M] ^ (y >> 1) ^ (-(int32_t)(y & 1) & MATRIX_A);
    }
    for (; i < N - 1; i++) {
      y = (key[i] & UPPER) | (key[i + 1] & LOWER);
      key[i] = key[i + (M - N)] ^ (y >> 1) ^ (-(int32_t)(y & 1) & MATRIX_A);
    }
    y = (key[N - 1] & UPPER) | (key[0] & LOWER);
    key[N - 1] = key[M - 1] ^ (y >> 1) ^ (-(int32_t)(y & 1) & MATRIX_A);
    pos = 0;
  }
  uint32_t next32() {
    if (pos == 624) refill();
    uint32_t y = key[pos++];
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return y;
  }
  double sample() {
    const int32_t a = (int32_t)(next32() >> 5), b = (int32_t)(next32() >> 6);
    return (a * 67108864.0 + b) / 9007199254740992.0;
  }
  double uniform(double low, double high) {
    const double range = high - low;
    return low + range * sample();
  }
  double legacy_gauss() {
    if (has_gauss) {
      const double t = gauss;
      has_gauss = 0;
      gauss = 0.0;
      return t;
    }
    double f, x1, x2, r2;
    do {
      x1 = 2.0 * sample() - 1.0;
      x2 = 2.0 * sample() - 1.0;
      r2 = x1 * x1 + x2 * x2;
    } while (r2 >= 1.0 || r2 == 0.0);
    f = std::sqrt(-2.0 * std::log(r2) / r2);
    gauss = f * x1;
    has_gauss = 1;
    return f * x2;
  }
  double normal(double loc, double scale) { return loc + scale * legacy_gauss(); }
  // RandomState.randint(low, high) for a range below 2^32: masked rejection, one 32-bit output per try
  int64_t randint(int64_t low, int64_t high) {
    const uint64_t rng = (uint64_t)(high - 1 - low);
    if (rng == 0) return low;
    uint32_t mask = (uint32_t)rng;
    mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
    uint32_t v;
    do v = next32() & mask; while (v > (uint32_t)rng);
    return low + (int64_t)v;
  }
};

constexpr int kMaxHi = 32;  // components per side the arrays below are sized for

// one slot's value at a border pixel, operation order of utils/gauss.py:77-88
struct SlotRot {
  double amp, cs, sn, ra, rb, wx, wy;
};
inline double slot_at(const SlotRot& s, double r, double c) {
  const double ua = r * s.cs - c * s.sn, ub = r * s.sn + c * s.cs;
  const double a = (s.ra - ua) / s.wy, b = (s.rb - ub) / s.wx;
  return s.amp * std::exp(-(a * a + b * b) / 2.0);
}

// upper bound of exp(-q / 2) for a slot's closest approach q to a border line; most slots are far from
// most lines, and there the bound need not be evaluated: exp(-40) < 4.25e-18
inline double line_bound(double q) {
  const double e = 0.5 * q * 0.999999;
  return e > 40.0 ? 4.25e-18 : std::exp(-e);
}

// jet.py:149-157: is the maximum of the summed components over the 4(N-1) border pixels
// below 0.01?  The reference evaluates every border pixel; here two rigorous shortcuts come
// first and the full evaluation only runs when neither decides:
//   (1) upper bound: a slot's maximum over a whole border LINE is amp * exp(-q/2) with q the
//       smallest value of its quadratic form on that line; the sum of the slots' largest line
//       values bounds every border pixel;
//   (2) lower bound: the exact sum (reference operation order) at the border pixel nearest to
//       where the strongest slot peaks on its best line.
// counts[0..2]: images decided by (1), by (2), by the full evaluation.
inline bool border_below(const SlotRot* s, int slots, int size, long long* counts) {
  const double thr = 0.01, last = (double)(size - 1);
  // Per border line (r = 0, r = N-1, c = 0, c = N-1) and slot: the slot's largest value on the whole
  // line, where along the line it sits and how fast it falls off there:
  //   value(t) = amp * exp(-(qmin + curv * (t - t0)^2) / 2),  t = the pixel's coordinate along the line
  struct LineSlot { double vmax, t0, curv; };
  LineSlot ls[4][2 * kMaxHi];
  double upper[4] = {0.0, 0.0, 0.0, 0.0}, best = -1.0;
  double best_r = 0.0, best_c = 0.0;
  bool bounded = slots <= 2 * kMaxHi;
  for (int k = 0; k < slots && bounded; ++k) {
    if (s[k].amp == 0.0) {
      for (int l = 0; l < 4; ++l) ls[l][k] = LineSlot{0.0, 0.0, 0.0};
      continue;
    }
    const double iy = 1.0 / (s[k].wy * s[k].wy), ix = 1.0 / (s[k].wx * s[k].wx);
    const double mrr = s[k].cs * s[k].cs * iy + s[k].sn * s[k].sn * ix;
    const double mcc = s[k].sn * s[k].sn * iy + s[k].cs * s[k].cs * ix;
    const double mrc = s[k].cs * s[k].sn * (ix - iy);
    const double det = mrr * mcc - mrc * mrc;
    // centre in (row, column): ra = cy cs - cx sn, rb = cy sn + cx cs  ->  cy, cx
    const double cy = s[k].ra * s[k].cs + s[k].rb * s[k].sn, cx = s[k].rb * s[k].cs - s[k].ra * s[k].sn;
    if (!(det > 0.0) || !std::isfinite(det) || !std::isfinite(cy) || !std::isfinite(cx) || !(mrr > 0.0) ||
        !(mcc > 0.0)) {
      bounded = false;
      break;
    }
    // (bounds only: one reciprocal per axis instead of a division per line; the rounding this adds is
    // ten orders of magnitude inside the 0.999999 factors)
    const double rows[2] = {0.0, last};
    const double imcc = 1.0 / mcc, imrr = 1.0 / mrr, aamp = std::fabs(s[k].amp);
    for (int e = 0; e < 2; ++e) {  // lines r = 0, r = N-1: along the line the column varies
      const double dr = rows[e] - cy, q = dr * dr * det * imcc;
      const double v = aamp * line_bound(q);
      const double t0 = cx - mrc * dr * imcc;
      ls[e][k] = LineSlot{v, t0, mcc * 0.999999};
      upper[e] += v;
      if (v > best) { best = v; best_r = rows[e]; best_c = t0; }
    }
    for (int e = 0; e < 2; ++e) {  // lines c = 0, c = N-1: the row varies
      const double dc = rows[e] - cx, q = dc * dc * det * imrr;
      const double v = aamp * line_bound(q);
      const double t0 = cy - mrc * dc * imrr;
      ls[2 + e][k] = LineSlot{v, t0, mrr * 0.999999};
      upper[2 + e] += v;
      if (v > best) { best = v; best_r = t0; best_c = rows[e]; }
    }
  }
  if (bounded) {
    const double u = std::fmax(std::fmax(upper[0], upper[1]), std::fmax(upper[2], upper[3]));
    if (u * 1.000001 < 0.0099) {  // (1): no border pixel can reach the threshold
      counts[0]++;
      return true;
    }
  }
  auto pixel = [&](double r, double c) {
    double sum = 0.0;
    for (int k = 0; k < slots; ++k) sum += (s[k].amp == 0.0) ? 0.0 : slot_at(s[k], r, c);
    return sum;
  };
  if (bounded && best >= 0.0) {
    const double r = std::fmin(std::fmax(std::nearbyint(best_r), 0.0), last);
    const double c = std::fmin(std::fmax(std::nearbyint(best_c), 0.0), last);
    if (!(pixel(r, c) < thr)) {  // (2) (a NaN sum makes the reference's `max < 0.01` false as well)
      counts[1]++;
      return false;
    }
  }
  counts[2]++;
  if (bounded) {
    // (3a) the full evaluation without the terms that cannot matter: on every line a slot only counts
    // where its own value can exceed `tiny`; what is left out is below slots * tiny per pixel.  The
    // verdict is taken from these sums unless the maximum lands within that margin of the threshold.
    const double tiny = 1e-14, margin = 2.0 * kMaxHi * tiny + 1e-15;
    bool all_amp_nonneg = true;
    for (int k = 0; k < slots; ++k) all_amp_nonneg = all_amp_nonneg && !(s[k].amp < 0.0);
    if (all_amp_nonneg) {
      double m = 0.0;
      bool above = false;
      for (int l = 0; l < 4 && !above; ++l) {
        int lo[2 * kMaxHi], hi[2 * kMaxHi], first = size, end = -1;
        for (int k = 0; k < slots; ++k) {
          lo[k] = 1; hi[k] = 0;
          const LineSlot& q = ls[l][k];
          if (!(q.vmax > tiny)) continue;
          // vmax * exp(-curv (t - t0)^2 / 2) > tiny  <=>  |t - t0| < sqrt(2 ln(vmax / tiny) / curv)
          const double reach = std::sqrt(2.0 * std::log(q.vmax / tiny) / q.curv) + 1.0;
          const double a0 = std::fmax(std::floor(q.t0 - reach), 0.0), a1 = std::fmin(std::ceil(q.t0 + reach), last);
          if (!(a0 <= a1)) continue;
          lo[k] = (int)a0; hi[k] = (int)a1;
          if (lo[k] < first) first = lo[k];
          if (hi[k] > end) end = hi[k];
        }
        const double fixed = (l & 1) ? last : 0.0;
        for (int i = first; i <= end; ++i) {
          double sum = 0.0;
          const double t = (double)i;
          for (int k = 0; k < slots; ++k)
            if (lo[k] <= i && i <= hi[k]) sum += (l < 2) ? slot_at(s[k], fixed, t) : slot_at(s[k], t, fixed);
          if (sum > m) m = sum;
          if (sum >= thr + margin) { above = true; break; }
        }
      }
      if (above || m >= thr + margin) return false;
      if (m + margin < thr) return true;
    }
  }
  // (3b) the reference's evaluation, term for term (np.max propagates NaN, and `nan < 0.01` is False)
  double m = -INFINITY;
  bool nan = false;
  for (int i = 0; i < size; ++i) {
    const double d = (double)i;
    const double v[4] = {pixel(0.0, d), pixel(last, d), pixel(d, 0.0), pixel(d, last)};
    for (int j = 0; j < 4; ++j) {
      if (v[j] != v[j]) nan = true;
      if (v[j] > m) m = v[j];
    }
  }
  return !nan && m < thr;
}

struct SqrtTable {
  double v[kMaxHi + 2];
  SqrtTable() {
    for (int i = 0; i < kMaxHi + 2; ++i) v[i] = std::sqrt((double)i);
  }
};

// One image of create_jet's chain (jet.py:54-125 and the draw at :168) on generator g.
// p [2*hi-1, 6], ex [2*hi-1, 2].
inline void pack_one_jet(Legacy& g, int size, int comp_lo, int hi, double* p, double* ex, int32_t* zoom_pad,
                         double* zoom_out, long long* counts) {
  const int slots = 2 * hi - 1;
  const double pi = 3.141592653589793, N = (double)size;
  const int mid = size / 2;
  double amp[kMaxHi] = {0}, x[kMaxHi] = {0}, y[kMaxHi] = {0}, sx[kMaxHi] = {0}, sy[kMaxHi] = {0},
         rot[kMaxHi] = {0}, beta[kMaxHi] = {0};
  static const SqrtTable roots;  // sqrt(0 .. kMaxHi), the same correctly rounded values std::sqrt returns
  const double* root = roots.v;
  const int k = (int)g.randint(comp_lo, (int64_t)hi + 1);
  rot[0] = g.uniform(0.0, pi);
  const double beta1 = g.uniform(0.0, 1.0);
  if (hi > 1) beta[1] = beta1;
  double bend = g.uniform(0.0, pi);
  const double tilt = g.uniform(0.0, pi / 2);
  const double sin_tilt = std::sin(tilt);
  for (int i = 0; i < k; ++i) {
    amp[i] = std::exp(-root[i] * g.normal(1.3, 0.2));
    if (i >= 1 && g.sample() < 0.1) amp[i] = 0.0;
    if (i >= 2) beta[i] = beta[1] * std::exp(-root[i - 1] * g.normal(0.5, 0.1));
    bend += g.normal(0.0, pi / 24);
    double cb, sb;
    sincos(bend, &sb, &cb);
    const double acb = std::fabs(cb), asb = std::fabs(sb);
    const double fill = acb < asb ? 1.0 / asb : (acb > asb ? 1.0 / acb : std::sqrt(2.0));
    const double dist = (double)i / (double)(k - 1) * N / 2 * fill * sin_tilt * 0.7;
    x[i] = dist * cb + mid;
    y[i] = dist * sb + mid;
    const double num = N / (double)k * fill * root[i + 1];
    const double p0 = num / g.uniform(3.0, 8.0), p1 = num / g.uniform(3.0, 8.0);
    if (p0 >= p1) { sx[i] = p0; sy[i] = p1; } else { sx[i] = p1; sy[i] = p0; }
    if (i >= 1) rot[i] = rot[i - 1] + g.normal(0.0, pi / 18);
  }
  const double mu = std::cos(tilt);
  const double off_x = g.uniform(-N / 20, N / 20), off_y = g.uniform(-N / 20, N / 20);
  std::memset(p, 0, sizeof(double) * slots * 6);
  std::memset(ex, 0, sizeof(double) * slots * 2);
  double away[kMaxHi];
  for (int i = 0; i < hi; ++i) {
    double toward = 1.0;  // (beta = 0: the expressions below give exactly 1)
    away[i] = 1.0;
    if (beta[i] != 0.0) {
      const double lorentz = 1.0 / std::sqrt(1.0 - beta[i] * beta[i]);
      toward = 1.0 / (lorentz * (1.0 - beta[i] * mu));
      away[i] = 1.0 / (lorentz * (1.0 + beta[i] * mu));
    }
    p[i * 6 + 0] = amp[i] * toward;
    p[i * 6 + 1] = x[i] + off_x;
    p[i * 6 + 2] = y[i] + off_y;
    p[i * 6 + 3] = sx[i];
    p[i * 6 + 4] = sy[i];
    p[i * 6 + 5] = rot[i];
    ex[i * 2 + 1] = beta[i];
  }
  for (int s = 0; s < slots; ++s) ex[s * 2] = tilt;
  if (g.sample() < 0.3) {  // counter jet: mirror slots 1.. through the image centre
    for (int i = 1; i < hi; ++i) {
      double* q = p + (size_t)(hi + i - 1) * 6;
      q[0] = amp[i] * away[i];
      q[1] = N - x[i] + off_x;
      q[2] = N - y[i] + off_y;
      q[3] = sx[i];
      q[4] = sy[i];
      q[5] = rot[i];
      ex[(hi + i - 1) * 2 + 1] = beta[i];
    }
  }
  SlotRot sr[2 * kMaxHi];
  for (int s = 0; s < slots; ++s) {
    const double* q = p + (size_t)s * 6;
    sr[s].amp = q[0];
    if (q[0] == 0.0) continue;
    const int twin = s - hi + 1;  // a counter-jet slot has the rotation of its forward twin
    if (s >= hi && sr[twin].amp != 0.0) {
      sr[s].sn = sr[twin].sn;
      sr[s].cs = sr[twin].cs;
    } else {
      sincos(q[5], &sr[s].sn, &sr[s].cs);
    }
    sr[s].ra = q[2] * sr[s].cs - q[1] * sr[s].sn;
    sr[s].rb = q[2] * sr[s].sn + q[1] * sr[s].cs;
    sr[s].wx = q[3];
    sr[s].wy = q[4];
  }
  *zoom_pad = -1;
  *zoom_out = 1.0;
  if (border_below(sr, slots, size, counts)) {
    const double zo = g.uniform(0.5, 1.0);
    *zoom_out = zo;
    *zoom_pad = (int32_t)(int64_t)((1.0 / zo - 1.0) * N / 2);
  }
}

}  // namespace pack
}  // namespace rsb

extern "C" int rsb_jet_pack(uint32_t* mt_key, int32_t* mt_pos, int32_t* has_gauss, double* cached_gauss,
                            int n, int size, int comp_lo, int comp_hi, double* params, double* extra,
                            int32_t* zoom_pad, double* zoom_out, long long* border_counts) {
  using namespace rsb::pack;
  RSB_ARG(mt_key && mt_pos && has_gauss && cached_gauss, "rsb_jet_pack: NULL generator state");
  RSB_ARG(params && extra && zoom_pad && zoom_out, "rsb_jet_pack: NULL output");
  RSB_ARG(n >= 0 && size >= 2, "rsb_jet_pack: n must be >= 0 and size >= 2 (got %d, %d)", n, size);
  RSB_ARG(comp_lo >= 1 && comp_lo <= comp_hi && comp_hi <= kMaxHi,
          "rsb_jet_pack: components must satisfy 1 <= lo <= hi <= %d (got %d, %d)", kMaxHi, comp_lo, comp_hi);
  RSB_ARG(*mt_pos >= 0 && *mt_pos <= 624, "rsb_jet_pack: generator position %d out of range", *mt_pos);
  Legacy g{mt_key, *mt_pos, *has_gauss, *cached_gauss};
  long long counts_local[3] = {0, 0, 0};
  long long* counts = border_counts ? border_counts : counts_local;
  const int slots = 2 * comp_hi - 1;
  for (int b = 0; b < n; ++b)
    pack_one_jet(g, size, comp_lo, comp_hi, params + (size_t)b * slots * 6, extra + (size_t)b * slots * 2,
                 zoom_pad + b, zoom_out + b, counts);
  *mt_pos = g.pos;
  *has_gauss = g.has_gauss;
  *cached_gauss = g.gauss;
  return 0;
}

// The source loop of create_survey (reference src/radiosim/survey/survey.py:49-107) on the same
// legacy stream: class by uniform(0, sum dist); class 0 draws a sub-jet (size, position, then the
// whole create_jet chain on size x size, border test included); classes 1 and 2 draw a wide or
// a narrow Gaussian.  Outputs are sized for the worst case (every source of one class):
//   src [n*num_sources, 6] (amp, x, y, sx, sy, rot): per image the wide rows, then the narrow
//   ones; src_offs [n, 3] = first wide row, first narrow row, end;
//   jet_* [n*num_sources]: sub-jets in draw order (parameters [.., 2*comp_hi-1, 6], zoom pad,
//   size, image index, row and column of the paste's corner: floor(pos - size/2)).
extern "C" int rsb_survey_pack(uint32_t* mt_key, int32_t* mt_pos, int32_t* has_gauss, double* cached_gauss,
                               int n, int size, int num_sources, const double* class_dist, int scale_sources,
                               int comp_lo, int comp_hi, double* src, int32_t* src_offs, double* jet_params,
                               int32_t* jet_zoom_pad, int32_t* jet_size, int32_t* jet_image,
                               int32_t* jet_row0, int32_t* jet_col0, int32_t* njets_out) {
  using namespace rsb::pack;
  RSB_ARG(mt_key && mt_pos && has_gauss && cached_gauss, "rsb_survey_pack: NULL generator state");
  RSB_ARG(class_dist && src && src_offs && jet_params && jet_zoom_pad && jet_size && jet_image && jet_row0 &&
              jet_col0 && njets_out, "rsb_survey_pack: NULL buffer");
  RSB_ARG(n >= 0 && size >= 2 && num_sources >= 0, "rsb_survey_pack: bad n / size / num_sources (%d, %d, %d)",
          n, size, num_sources);
  RSB_ARG(comp_lo >= 2 && comp_lo <= comp_hi && comp_hi <= kMaxHi,
          "rsb_survey_pack: components must satisfy 2 <= lo <= hi <= %d (got %d, %d)", kMaxHi, comp_lo, comp_hi);
  RSB_ARG(*mt_pos >= 0 && *mt_pos <= 624, "rsb_survey_pack: generator position %d out of range", *mt_pos);
  Legacy g{mt_key, *mt_pos, *has_gauss, *cached_gauss};
  long long counts[3] = {0, 0, 0};
  const int slots = 2 * comp_hi - 1;
  const double pi = 3.141592653589793, N = (double)size;
  const double total = class_dist[0] + class_dist[1] + class_dist[2];
  const double c01 = class_dist[0] + class_dist[1];
  double* narrow = new (std::nothrow) double[(size_t)(num_sources > 0 ? num_sources : 1) * 6];
  RSB_ARG(narrow != nullptr, "rsb_survey_pack: out of host memory");
  double extra_unused[2 * (2 * kMaxHi)];
  double zoom_out_unused;
  int nj = 0;
  int64_t row = 0;
  for (int b = 0; b < n; ++b) {
    const int64_t first = row;
    int nnarrow = 0;
    for (int q = 0; q < num_sources; ++q) {
      const double u = g.uniform(0.0, total);
      if (u < class_dist[0]) {
        const int64_t lo = scale_sources ? (int64_t)(N / 10.24) : 100, hi = scale_sources ? (int64_t)(N / 5.12) : 200;
        const int js = (int)g.randint(lo, hi);
        const double pa = g.sample() * N, pb = g.sample() * N;
        if (js < 2) {
          delete[] narrow;
          RSB_ARG(false, "rsb_survey_pack: sub-jet size %d too small (image size %d)", js, size);
        }
        pack_one_jet(g, js, comp_lo, comp_hi, jet_params + (size_t)nj * slots * 6, extra_unused,
                     jet_zoom_pad + nj, &zoom_out_unused, counts);
        jet_size[nj] = js;
        jet_image[nj] = b;
        jet_row0[nj] = (int32_t)std::floor(pa - (double)js / 2);
        jet_col0[nj] = (int32_t)std::floor(pb - (double)js / 2);
        ++nj;
      } else {
        const bool wide = u < c01;
        double sx, sy;
        if (wide) {
          if (scale_sources) {
            sx = N * g.uniform(1 / 40.96, 1 / 20.48);
            sy = sx + g.normal(0.0, N / 204.8);
          } else {
            sx = g.uniform(25.0, 50.0);
            sy = sx + g.normal(0.0, 5.0);
          }
        } else {
          if (scale_sources) {
            sx = N * g.uniform(1 / 204.8, 1 / 102.4);
            sy = sx + g.normal(0.0, N / 1024);
          } else {
            sx = g.uniform(5.0, 10.0);
            sy = sx + g.normal(0.0, 1.0);
          }
        }
        const double px = g.sample() * N, py = g.sample() * N;
        const double amp = g.sample();
        const double rot = g.uniform(0.0, pi);
        double* d = wide ? src + (size_t)(row++) * 6 : narrow + (size_t)(nnarrow++) * 6;
        d[0] = amp; d[1] = px; d[2] = py; d[3] = sx; d[4] = sy; d[5] = rot;
      }
    }
    const int64_t mid = row;
    std::memcpy(src + (size_t)row * 6, narrow, sizeof(double) * 6 * (size_t)nnarrow);
    row += nnarrow;
    src_offs[b * 3] = (int32_t)first;
    src_offs[b * 3 + 1] = (int32_t)mid;
    src_offs[b * 3 + 2] = (int32_t)row;
  }
  delete[] narrow;
  *njets_out = nj;
  *mt_pos = g.pos;
  *has_gauss = g.has_gauss;
  *cached_gauss = g.gauss;
  return 0;
}


// `count` draws of RandomState.normal() (loc 0, scale 1: the value is legacy_gauss itself) as
// float32 -- the noise fields of add_noise, reference src/radiosim/utils/utils.py:277,291
// (np.random.normal(size=img_shape)), drawn in stream order.  Same state hand-over as rsb_jet_pack.
// The stream itself is sequential, but only its cheapest part is: the calling thread runs MT19937 and
// tempers its words, a block of candidate pairs at a time (four words per candidate, accepted or not, so
// candidate k of a block sits at a known place); worker threads then do everything else -- the polar
// method's rejection test, sqrt(-2 log(r2) / r2) * {x2, x1} for the accepted pairs -- on their share of the
// candidates and write the results where a prefix sum of the accepted counts puts them.  Every value goes
// through the same double operations in the same order as in legacy_gauss(), so the output, the cached
// second variate and the stream position equal one call after the other.
namespace rsb {
namespace pack {
inline int fill_threads() {
  if (const char* e = getenv("RSB_HOST_THREADS")) {
    const int v = atoi(e);
    if (v >= 1) return v > 64 ? 64 : v;
  }
  unsigned hw = std::thread::hardware_concurrency();
  if (const char* e = getenv("LOCAL_WORLD_SIZE")) {  // one process per GPU (torchrun): the cores are shared
    const int ranks = atoi(e);
    if (ranks > 1) hw /= (unsigned)ranks;
  }
  return hw >= 16 ? 8 : (hw >= 4 ? (int)hw / 2 : 1);
}
// n tempered words of the stream into w (the generator ends up behind them)
inline void gen_words(Legacy& g, uint32_t* w, long long n) {
  while (n > 0) {
    if (g.pos == 624) g.refill();
    const int step = (int)((624 - g.pos) < n ? (624 - g.pos) : n);
    const uint32_t* k = g.key + g.pos;
    for (int q = 0; q < step; ++q) {
      uint32_t y = k[q];
      y ^= (y >> 11);
      y ^= (y << 7) & 0x9d2c5680u;
      y ^= (y << 15) & 0xefc60000u;
      y ^= (y >> 18);
      w[q] = y;
    }
    g.pos += step;
    w += step;
    n -= step;
  }
}
inline void skip_words(Legacy& g, long long n) {
  while (n > 0) {
    if (g.pos == 624) g.refill();
    const int step = (int)((624 - g.pos) < n ? (624 - g.pos) : n);
    g.pos += step;
    n -= step;
  }
}
}  // namespace pack
}  // namespace rsb

extern "C" int rsb_legacy_normal_fill(uint32_t* mt_key, int32_t* mt_pos, int32_t* has_gauss,
                                      double* cached_gauss, long long count, float* out) {
  using namespace rsb::pack;
  RSB_ARG(mt_key && mt_pos && has_gauss && cached_gauss && (out || count == 0),
          "rsb_legacy_normal_fill: NULL argument");
  RSB_ARG(count >= 0 && *mt_pos >= 0 && *mt_pos <= 624, "rsb_legacy_normal_fill: bad count / generator position");
  Legacy g{mt_key, *mt_pos, *has_gauss, *cached_gauss};
  long long i = 0;
  if (count > 0 && g.has_gauss) out[i++] = (float)g.legacy_gauss();  // the value a previous call left behind
  const long long pairs = (count - i) / 2;
  const int threads = fill_threads();
  constexpr long long kSmall = 32768;  // pairs below which threads do not pay
  if (pairs >= kSmall && threads > 1) {
    constexpr long long kCand = 1 << 19;  // candidates per block (8 MB of words)
    try {
      // scratch kept between calls (add_noise fills three fields per bundle): ~22 MB at most
      static thread_local std::vector<uint32_t> words;
      static thread_local std::vector<double> X1, X2, R2;
      static thread_local std::vector<int32_t> cidx;
      {
        long long cap = (long long)((double)pairs * 1.2732) + 64;
        if (cap > kCand) cap = kCand;
        if ((long long)X1.size() < cap) {
          words.resize((size_t)4 * cap); X1.resize((size_t)cap); X2.resize((size_t)cap); R2.resize((size_t)cap);
          cidx.resize((size_t)cap);
        }
      }
      std::vector<long long> cnt((size_t)threads);
      long long done = 0;
      double under = 0.985;
      if (const char* e = getenv("RSB_DEBUG_FILL_FACTOR")) under = atof(e);  // (tests: > 1 forces the overshoot path)
      while (pairs - done >= kSmall) {
        const long long need = pairs - done;
        // a little under the expected requirement (acceptance pi/4), so that the last block rarely overshoots
        long long nc = (long long)((double)need * 1.2732 * under);
        if (nc > kCand) nc = kCand;
        uint32_t snap[624];
        std::memcpy(snap, g.key, sizeof(snap));
        const int snap_pos = g.pos;
        gen_words(g, words.data(), 4 * nc);
        std::atomic<int> arrived{0};
        const long long per = (nc + threads - 1) / threads;
        float* dst = out + i + 2 * done;
        std::vector<std::thread> pool;
        // (plain pointers: a lambda that names a thread_local vector would see the WORKER's own, empty one)
        uint32_t* const Wp = words.data();
        double* const pX1 = X1.data();
        double* const pX2 = X2.data();
        double* const pR2 = R2.data();
        int32_t* const pC = cidx.data();
        for (int t = 0; t < threads; ++t) {
          pool.emplace_back([&, t] {
            const long long a = t * per, e = (a + per) < nc ? (a + per) : nc;
            long long k = 0;
            for (long long c = a; c < e; ++c) {
              const uint32_t* w = Wp + (size_t)4 * c;
              const double s1 = ((double)(int32_t)(w[0] >> 5) * 67108864.0 + (double)(int32_t)(w[1] >> 6)) / 9007199254740992.0;
              const double s2 = ((double)(int32_t)(w[2] >> 5) * 67108864.0 + (double)(int32_t)(w[3] >> 6)) / 9007199254740992.0;
              const double x1 = 2.0 * s1 - 1.0, x2 = 2.0 * s2 - 1.0, r2 = x1 * x1 + x2 * x2;
              pX1[a + k] = x1; pX2[a + k] = x2; pR2[a + k] = r2; pC[a + k] = (int32_t)c;
              k += (r2 >= 1.0 || r2 == 0.0) ? 0 : 1;   // rejected: overwritten by the next candidate
            }
            cnt[t] = a < e ? k : 0;
            arrived.fetch_add(1, std::memory_order_acq_rel);
            while (arrived.load(std::memory_order_acquire) < threads) std::this_thread::yield();
            long long off = 0;
            for (int u = 0; u < t; ++u) off += cnt[u];
            long long mine = cnt[t];
            if (off + mine > need) mine = need > off ? need - off : 0;  // (the last block may overshoot)
            for (long long q = 0; q < mine; ++q) {
              const double r2 = pR2[a + q];
              const double f = std::sqrt(-2.0 * std::log(r2) / r2);
              dst[2 * (off + q)] = (float)(f * pX2[a + q]);      // returned first
              dst[2 * (off + q) + 1] = (float)(f * pX1[a + q]);  // the cached second variate
            }
          });
        }
        for (auto& th : pool) th.join();
        long long total = 0;
        for (int t = 0; t < threads; ++t) total += cnt[t];
        if (total > need) {
          // overshoot: the stream must stop behind the candidate that gave the last pair needed
          long long off = 0, last_cand = -1;
          for (int t = 0; t < threads && last_cand < 0; ++t) {
            if (off + cnt[t] >= need) last_cand = pC[(size_t)(t * per + (need - off - 1))];
            off += cnt[t];
          }
          std::memcpy(g.key, snap, sizeof(snap));
          g.pos = snap_pos;
          skip_words(g, 4 * (last_cand + 1));
          total = need;
        }
        done += total;
      }
      i += 2 * done;
    } catch (...) {
      rsb::set_error("rsb_legacy_normal_fill: out of host memory or threads");
      return 1;
    }
  }
  for (; i < count; ++i) out[i] = (float)g.legacy_gauss();  // the rest (and an odd count's cached variate)
  *mt_pos = g.pos;
  *has_gauss = g.has_gauss;
  *cached_gauss = g.gauss;
  return 0;
}


// ---- ppdisk: the RNG chain of generate_proto_set / create_proto ---------------------------
// reference src/radiosim/ppdisks/ppdisks.py: per image alpha = randint(*alpha_range) (:74), ratio =
// randint(*ratio_range) / randint(*ratio_range) (:75), size_ratio = uniform(*size_ratio_range) (:76);
// then create_proto: choice of one of four brightness fits (:474), core a = randint(3, 15) (:305),
// sc = randint(5, 10) (:135), core scale = exp(m d + n) + normal(0, 0.05 exp(..)) (:477-483),
// num_circ = floor(normal(3, 0.5)) (:137), per ring randint(10, 35), randint(5, 10), its scale
// (:360-361, :483); crop and resize target (:154-167).  Same legacy stream and state hand-over as
// rsb_jet_pack.  cos_tab / sin_tab [alpha_hi]: float32(cos / sin(deg2rad(alpha))) as numpy computes
// them (the masks are bit-exact functions of these two numbers, so they come from the caller).
// comps [comp_cap] 48-byte and images [n] 32-byte records as rsb_ppdisk_render takes them.
namespace rsb {
namespace pack {
struct PpComp { float ca, sa, a2_in, b2_in, a2_out, b2_out, scale; int sigma, image, pad[3]; };
struct PpImage { int comp0, ncomp, lo, hi, new_size, pad[3]; };
static_assert(sizeof(PpComp) == 48 && sizeof(PpImage) == 32, "ppdisk record layout");
inline int pp_smooth_factor(double a) { return a < 15 ? 2 : a < 20 ? 4 : a < 30 ? 6 : a < 40 ? 9 : 12; }
}  // namespace pack
}  // namespace rsb

extern "C" int rsb_ppdisk_pack(uint32_t* mt_key, int32_t* mt_pos, int32_t* has_gauss, double* cached_gauss,
                               int n, int img_size, int alpha_lo, int alpha_hi, int ratio_lo, int ratio_hi,
                               double size_ratio_lo, double size_ratio_hi, const float* cos_tab,
                               const float* sin_tab, void* comps, int comp_cap, void* images,
                               int32_t* n_comps_out) {
  using namespace rsb::pack;
  RSB_ARG(mt_key && mt_pos && has_gauss && cached_gauss && cos_tab && sin_tab && comps && images && n_comps_out,
          "rsb_ppdisk_pack: NULL argument");
  RSB_ARG(n >= 0 && img_size >= 2 && 0 <= alpha_lo && alpha_lo < alpha_hi && 0 < ratio_lo && ratio_lo < ratio_hi,
          "rsb_ppdisk_pack: bad ranges");
  RSB_ARG(*mt_pos >= 0 && *mt_pos <= 624, "rsb_ppdisk_pack: generator position %d out of range", *mt_pos);
  static const double fits[4][2] = {{-0.08908080536281453, -4.808388735775005},
                                    {-0.04053517757866287, -5.718461569072661},
                                    {-0.023989028595164514, -6.22105762345096},
                                    {-0.011198837787046627, -6.701733964123176}};
  Legacy g{mt_key, *mt_pos, *has_gauss, *cached_gauss};
  PpComp* C = (PpComp*)comps;
  PpImage* I = (PpImage*)images;
  int nc = 0;
  bool overflow = false;
  auto push = [&](float ca, float sa, double a2i, double b2i, double a2o, double b2o, double scale, int sigma, int img) {
    if (nc < comp_cap) {
      PpComp c;
      c.ca = ca; c.sa = sa; c.a2_in = (float)a2i; c.b2_in = (float)b2i; c.a2_out = (float)a2o; c.b2_out = (float)b2o;
      c.scale = (float)scale; c.sigma = sigma; c.image = img; c.pad[0] = c.pad[1] = c.pad[2] = 0;
      C[nc] = c;
    } else {
      overflow = true;
    }
    ++nc;
  };
  for (int b = 0; b < n; ++b) {
    const int alpha = (int)g.randint(alpha_lo, alpha_hi);
    const double r_num = (double)g.randint(ratio_lo, ratio_hi), r_den = (double)g.randint(ratio_lo, ratio_hi);
    const double ratio = r_num / r_den;
    const double size_ratio = g.uniform(size_ratio_lo, size_ratio_hi);
    const int fit = (int)g.randint(0, 4);
    const double m = fits[fit][0], nn = fits[fit][1];
    auto scale = [&](double d) {
      const double e = std::exp(m * d + nn);
      return e + g.normal(0.0, e * 0.05);
    };
    const float ca = cos_tab[alpha], sa = sin_tab[alpha];
    const double a = (double)g.randint(3, 15);
    const double bb = a / ratio;
    const double d_core = (a + bb) / 2;
    const int sc = (int)g.randint(5, 10);
    I[b].comp0 = nc;
    {
      const double s_core = scale(d_core);
      push(ca, sa, -1.0, 1.0, a * a, bb * bb, s_core, sc, b);
    }
    const int num_circ = (int)std::floor(g.normal(3.0, 0.5));
    double a_max = d_core;
    for (int i = 0; i < num_circ; ++i) {
      const double ra = (double)(g.randint(10, 35) + (i + 1) * 30);
      const double ex = (double)g.randint(5, 10);
      const double d_ring = ((ra + ex + (ra + ex) / ratio) / 2 + (ra + ra / ratio) / 2) / 2;
      const int s = pp_smooth_factor(ra);
      const double s_ring = scale(d_ring);
      const double rin = ra / ratio, rout = (ra + ex) / ratio;
      push(ca, sa, ra * ra, rin * rin, (ra + ex) * (ra + ex), rout * rout, s_ring, s, b);
      double extent = std::fmax(std::fmax(ra, ra + ex), std::fmax(rin, rout)) + 6 * s;
      a_max = std::fmax(extent, a_max);
    }
    const double half = 450.0;
    a_max = std::fmin(half, a_max);
    I[b].ncomp = nc - I[b].comp0;
    I[b].lo = (int)(half - a_max);
    I[b].hi = (int)(half + a_max);
    int nw = (int)(size_ratio * img_size);
    if (nw % 2) nw -= 1;
    I[b].new_size = nw;
    I[b].pad[0] = I[b].pad[1] = I[b].pad[2] = 0;
  }
  *n_comps_out = nc;
  *mt_pos = g.pos;
  *has_gauss = g.has_gauss;
  *cached_gauss = g.gauss;
  RSB_ARG(!overflow, "rsb_ppdisk_pack: %d components do not fit the buffer of %d", nc, comp_cap);
  return 0;
}
