// Native host-side recipe generator for the MOJAVE-like sources: the parameter draws of
// gen_compact / gen_one_jet / gen_two_jet (reference src/radiosim/mojave/mojave.py:264-662,
// helpers gen_shocks :124-219, add_swirl :222-261) replayed draw for draw on the CHILD generators
// that create_mojave spawns (:56-64), in C++ instead of Python.  Host code only.
//
// What is restated here (numpy 2.x, numpy/random/bit_generator.pyx, src/pcg64, src/distributions):
//   SeedSequence : child i of a parent is SeedSequence(entropy, spawn_key + (i,)); its pool comes
//                  from the 32-bit hashmix / mix rounds, generate_state(4 x uint64) seeds PCG64
//   PCG64        : 128-bit LCG (multiplier 0x2360ED051FC65DA44385DF649FCCF645), XSL-RR output;
//                  next_double = (next64 >> 11) / 2^53; next_uint32 hands out the two halves of a draw
//   normal / standard_exponential : the 256-layer ziggurats (tables read out of numpy's own binary,
//                  csrc/ziggurat_tables.h), uniform(l, h) = l + (h - l) * next_double,
//                  power(a) = (-expm1(-standard_exponential))^(1/a), integers(l, h) = Lemire's
//                  rejection on next_uint32 for ranges below 2^32
//   scipy.stats  : expon.rvs = standard_exponential * scale + loc; skewnorm.rvs = the two-normal
//                  construction of skewnorm_gen._rvs
// tests/test_mojave_native.py compares the streams with numpy itself and the recipes with the Python
// packer (radiosim_b200/mojave/mojave.py), which stays as the reference implementation of this file.
//
// One thing moves to the device: a shock's amplitude is "pixel of the jet so far x uniform draw"
// (mojave.py:190).  The draw is made here; the pixel -- which needs the max-normalised 1-D profiles of
// the blobs before it -- is evaluated by mojave::blob_kernel (csrc/mojave.cu), blob after blob.
#include <cmath>
#include <cstdint>
#include <cstring>

#include "common.cuh"
#include "ziggurat_tables.h"

namespace rsb {
namespace mjpack {

typedef unsigned __int128 u128;

struct Pcg64 {
  u128 state, inc;
  int has_uint32;
  uint32_t uinteger;

  static u128 mult() { return ((u128)0x2360ED051FC65DA4ULL << 64) | 0x4385DF649FCCF645ULL; }
  void step() { state = state * mult() + inc; }
  void seed(uint64_t s_hi, uint64_t s_lo, uint64_t i_hi, uint64_t i_lo) {
    const u128 initstate = ((u128)s_hi << 64) | s_lo, initseq = ((u128)i_hi << 64) | i_lo;
    state = 0;
    inc = (initseq << 1) | 1;
    step();
    state += initstate;
    step();
    has_uint32 = 0;
    uinteger = 0;
  }
  uint64_t next64() {
    step();
    const uint64_t hi = (uint64_t)(state >> 64), lo = (uint64_t)state;
    const uint64_t x = hi ^ lo;
    const unsigned rot = (unsigned)(state >> 122);
    return (x >> rot) | (x << ((64 - rot) & 63));
  }
  uint32_t next32() {
    if (has_uint32) {
      has_uint32 = 0;
      return uinteger;
    }
    const uint64_t n = next64();
    has_uint32 = 1;
    uinteger = (uint32_t)(n >> 32);
    return (uint32_t)(n & 0xffffffffu);
  }
  double next_double() { return (double)(next64() >> 11) * (1.0 / 9007199254740992.0); }

  double standard_normal() {
    for (;;) {
      uint64_t r = next64();
      const int idx = (int)(r & 0xff);
      r >>= 8;
      const int sign = (int)(r & 0x1);
      const uint64_t rabs = (r >> 1) & 0x000fffffffffffffULL;
      double x = (double)rabs * zig::wi_double[idx];
      if (sign & 0x1) x = -x;
      if (rabs < zig::ki_double[idx]) return x;
      if (idx == 0) {
        for (;;) {
          const double xx = -zig::kNorInvR * std::log1p(-next_double());
          const double yy = -std::log1p(-next_double());
          if (yy + yy > xx * xx) return ((rabs >> 8) & 0x1) ? -(zig::kNorR + xx) : zig::kNorR + xx;
        }
      } else {
        if (((zig::fi_double[idx - 1] - zig::fi_double[idx]) * next_double() + zig::fi_double[idx]) <
            std::exp(-0.5 * x * x))
          return x;
      }
    }
  }
  double standard_exponential() {
    for (;;) {
      uint64_t ri = next64();
      ri >>= 3;
      const int idx = (int)(ri & 0xFF);
      ri >>= 8;
      const double x = (double)ri * zig::we_double[idx];
      if (ri < zig::ke_double[idx]) return x;
      if (idx == 0) return zig::kExpR - std::log1p(-next_double());
      if ((zig::fe_double[idx - 1] - zig::fe_double[idx]) * next_double() + zig::fe_double[idx] < std::exp(-x))
        return x;
    }
  }
  double normal(double loc, double scale) { return loc + scale * standard_normal(); }
  double uniform(double low, double high) { return low + (high - low) * next_double(); }
  double power(double a) { return std::pow(-std::expm1(-standard_exponential()), 1.0 / a); }
  // Generator.integers(low, high) for high - low <= 2^32 (int64 path, endpoint False)
  int64_t integers(int64_t low, int64_t high) {
    const uint64_t rng = (uint64_t)(high - 1 - low);
    if (rng == 0) return low;
    if (rng == 0xFFFFFFFFULL) return low + (int64_t)next32();
    const uint32_t rng_excl = (uint32_t)rng + 1u;
    uint64_t m = (uint64_t)next32() * (uint64_t)rng_excl;
    uint32_t leftover = (uint32_t)(m & 0xFFFFFFFFULL);
    if (leftover < rng_excl) {
      const uint32_t threshold = (uint32_t)((0xFFFFFFFFULL - rng) % rng_excl);
      while (leftover < threshold) {
        m = (uint64_t)next32() * (uint64_t)rng_excl;
        leftover = (uint32_t)(m & 0xFFFFFFFFULL);
      }
    }
    return low + (int64_t)(m >> 32);
  }
  double expon(double loc, double scale) { return standard_exponential() * scale + loc; }
  double skewnorm(double a, double loc, double scale) {
    const double u0 = standard_normal();
    const double v = standard_normal();
    const double d = a / std::sqrt(1.0 + a * a);
    const double u1 = d * u0 + v * std::sqrt(1.0 - d * d);
    return (u0 >= 0.0 ? u1 : -u1) * scale + loc;
  }
};

// ---- SeedSequence (pool size 4) -------------------------------------------------------------
constexpr uint32_t kInitA = 0x43b0d7e5u, kMultA = 0x931e8875u, kInitB = 0x8b51f9ddu, kMultB = 0x58f38dedu;
constexpr uint32_t kMixL = 0xca01f9ddu, kMixR = 0x4973f715u;
constexpr int kPool = 4, kXshift = 16;

inline uint32_t hashmix(uint32_t value, uint32_t& hash_const) {
  value ^= hash_const;
  hash_const *= kMultA;
  value *= hash_const;
  value ^= value >> kXshift;
  return value;
}
inline uint32_t mix(uint32_t x, uint32_t y) {
  uint32_t r = kMixL * x - kMixR * y;
  r ^= r >> kXshift;
  return r;
}
// entropy words (already padded to the pool size when a spawn key follows) ++ spawn key words
inline void seed_pool(const uint32_t* words, int n, uint32_t pool[kPool]) {
  uint32_t hash_const = kInitA;
  for (int i = 0; i < kPool; ++i) pool[i] = hashmix(i < n ? words[i] : 0u, hash_const);
  for (int i_src = 0; i_src < kPool; ++i_src)
    for (int i_dst = 0; i_dst < kPool; ++i_dst)
      if (i_src != i_dst) pool[i_dst] = mix(pool[i_dst], hashmix(pool[i_src], hash_const));
  for (int i_src = kPool; i_src < n; ++i_src)
    for (int i_dst = 0; i_dst < kPool; ++i_dst) pool[i_dst] = mix(pool[i_dst], hashmix(words[i_src], hash_const));
}
inline void generate_state4(const uint32_t pool[kPool], uint64_t out[4]) {
  uint32_t hash_const = kInitB, w[8];
  for (int i = 0; i < 8; ++i) {
    uint32_t v = pool[i % kPool];
    v ^= hash_const;
    hash_const *= kMultB;
    v *= hash_const;
    v ^= v >> kXshift;
    w[i] = v;
  }
  for (int i = 0; i < 4; ++i) out[i] = (uint64_t)w[2 * i] | ((uint64_t)w[2 * i + 1] << 32);
}

constexpr int kMaxKey = 60;
struct Parent {
  uint32_t words[kMaxKey + 4];  // run entropy (padded) ++ spawn key; the child index goes behind
  int n;
};
inline Pcg64 child_generator(const Parent& p, uint32_t index) {
  uint32_t words[kMaxKey + 5];
  std::memcpy(words, p.words, sizeof(uint32_t) * p.n);
  words[p.n] = index;
  uint32_t pool[kPool];
  seed_pool(words, p.n + 1, pool);
  uint64_t s[4];
  generate_state4(pool, s);
  Pcg64 g;
  g.seed(s[0], s[1], s[2], s[3]);
  return g;
}

// ---- recipes ------------------------------------------------------------------------------
struct Gauss { double amp, x, y, sx, sy, rot; };
struct Recipe {  // = rsb::mojave::Recipe (csrc/mojave.cu), 224 bytes
  Gauss g[3];
  double strength[2], radius[2];
  double angle, amp;
  int blob0[2], nblob[2];
  int flip1, compact;
  int pad[2];
};
static_assert(sizeof(Recipe) == 224, "Recipe layout");
struct BlobParam {  // one skewed-Gaussian blob as drawn (gauss.py:100-150 arguments), 64 bytes
  double x, y;        // centre: column, row
  double amp;         // amplitude, or for a shock the factor its reference pixel is multiplied with
  double width, length, a;
  int ref_r, ref_c;   // shock: pixel of the jet so far that scales the amplitude (mojave.py:190); -1: none
  int pad[2];
};
static_assert(sizeof(BlobParam) == 64, "BlobParam layout");

struct Out {
  BlobParam* blobs;
  int nblobs, cap;
  bool overflow;
  void push(const BlobParam& b) {
    if (nblobs < cap) blobs[nblobs] = b;
    else overflow = true;
    ++nblobs;
  }
};

inline void sort2(double& lo, double& hi) { if (hi < lo) { const double t = lo; lo = hi; hi = t; } }

inline void base_jet(Out& o, int size, double amp, double width, double length, double a) {
  BlobParam b;
  b.x = size / 2.0; b.y = size / 2.0; b.amp = amp; b.width = width; b.length = length; b.a = a;
  b.ref_r = b.ref_c = -1; b.pad[0] = b.pad[1] = 0;
  o.push(b);
}
// gen_shocks :124-219
inline void add_shocks(Pcg64& g, Out& o, int size, int n_shock, double length, double width, double sx) {
  const double step = (2 * length - sx) / n_shock;
  double dist = step;
  for (int k = 0; k < n_shock; ++k) {
    if (g.uniform(0.0, 1.0) > 2.0 / 3.0) continue;
    const double px = size / 2.0 + dist, py = size / 2.0;
    const long long r = (long long)py, c = (long long)px;   // int(): truncation toward zero
    if (!(-size <= r && r < size && -size <= c && c < size)) continue;   // the reference's IndexError
    BlobParam b;
    b.amp = g.uniform(0.03, 0.6);
    double lo = 2.0, hi = width;
    sort2(lo, hi);
    b.width = g.uniform(lo, hi);
    lo = b.width; hi = length / n_shock;
    sort2(lo, hi);
    b.length = g.uniform(lo, hi);
    b.a = g.uniform(5.0, 9.0);
    b.x = (double)(long long)px;
    b.y = (double)(long long)py;
    b.ref_r = (int)(((r % size) + size) % size);
    b.ref_c = (int)(((c % size) + size) % size);
    b.pad[0] = b.pad[1] = 0;
    o.push(b);
    dist += step;
  }
}
inline void core_widths(Pcg64& g, double dmin, double dmax, double& sx, double& sy, double& rot) {
  sx = g.uniform(dmin, dmax) * 1.5;
  sy = g.uniform(dmin, dmax) * 1.5;
  while (sx / sy < 0.5 || sx / sy > 2.0) sy = g.uniform(dmin, dmax) * 1.5;
  rot = g.uniform(0.0, 2 * 3.141592653589793);
}
inline int shock_count(Pcg64& g) {
  // int(np.round(rng.power(0.5), decimals=1) * 10)
  return (int)(std::nearbyint(g.power(0.5) * 10.0) / 10.0 * 10.0);
}

inline void recipe_two_jet(Pcg64& g, int size, Recipe& R, Out& o) {
  double dev[5];
  for (int i = 0; i < 5; ++i) dev[i] = 1 + g.normal(0.0, 0.1);
  const double amp = g.expon(1.715e-4, 1.985e-2);
  const double r_w = g.skewnorm(4.37, 22.71, 16.85) / 10;
  const double r_l = g.skewnorm(8.62, 69.60, 54.73) / 4;
  const double r_amp = amp * g.power(0.7);
  const double l_w = r_w * dev[2], l_l = r_l * dev[1], l_amp = r_amp * dev[0];
  const double d[4] = {r_w / 2, r_l / 5, l_w / 2, l_l / 5};
  double dmin = d[0], dmax = d[0];
  for (int i = 1; i < 4; ++i) { dmin = std::fmin(dmin, d[i]); dmax = std::fmax(dmax, d[i]); }
  double sx, sy, rot;
  core_widths(g, dmin, dmax, sx, sy, rot);
  const int n_r = shock_count(g), n_l = shock_count(g);
  R.blob0[0] = o.nblobs;
  base_jet(o, size, r_amp, r_w, r_l, 4.0);
  if (n_r > 0) add_shocks(g, o, size, n_r, r_l, r_w, sx);
  R.nblob[0] = o.nblobs - R.blob0[0];
  // (the reference builds both jets first and draws the right jet's shocks before the left jet's:
  //  the blobs of the left jet follow, its shocks are drawn next)
  R.blob0[1] = o.nblobs;
  base_jet(o, size, l_amp, l_w, l_l, 4.0);
  if (n_l > 0) add_shocks(g, o, size, n_l, l_l, l_w, sx);
  R.nblob[1] = o.nblobs - R.blob0[1];
  R.strength[0] = g.normal(0.0, 0.2);
  R.radius[0] = g.uniform(50.0, 200.0);
  const double d0 = 1 + g.normal(0.0, 0.1), d1 = 1 + g.normal(0.0, 0.1);
  R.strength[1] = R.strength[0] * d0;
  R.radius[1] = R.radius[0] * d1;
  R.flip1 = 1;
  R.g[0] = Gauss{amp, size / 2.0, size / 2.0, sx, sy, rot};
  R.angle = g.uniform(0.0, 360.0);
}

inline void recipe_one_jet(Pcg64& g, int size, Recipe& R, Out& o) {
  const double amp = g.expon(1.715e-4, 1.985e-2);
  const double width = g.skewnorm(4.37, 22.71, 16.85) / 12;
  const double length = g.skewnorm(8.62, 69.60, 54.73) / 6.5;
  const double jet_amp = amp * g.power(0.7);
  const double da = width / 2, db = length / 5;
  double sx, sy, rot;
  core_widths(g, std::fmin(da, db), std::fmax(da, db), sx, sy, rot);
  const int n_s = shock_count(g);
  R.blob0[0] = o.nblobs;
  base_jet(o, size, jet_amp, width, length, 4.0);
  if (n_s > 0) add_shocks(g, o, size, n_s, length, width, sx);
  R.nblob[0] = o.nblobs - R.blob0[0];
  R.strength[0] = g.normal(0.0, 0.2);
  R.radius[0] = g.uniform(50.0, 200.0);
  R.g[0] = Gauss{amp, size / 2.0, size / 2.0, sx, sy, rot};
  R.angle = g.uniform(0.0, 360.0);
}

inline double bounded_width(Pcg64& g, double lo, double hi) {
  double v = g.uniform(lo, hi) * 1.7;
  int tries = 0;
  while (v > 15) {
    if (tries > 10) { v /= 5; break; }
    v = g.uniform(lo, hi) * 1.7;
    ++tries;
  }
  return v;
}

inline void recipe_compact(Pcg64& g, int size, Recipe& R, Out& o) {
  const int ncomps = (int)g.integers(1, 4);
  const bool outflow = g.uniform(0.0, 1.0) > 0.5;
  const double amp = g.expon(5.54e-05, 0.02);
  const double jet_amp = amp * g.normal(2.0 / 3.0, 0.05);
  const double width = g.skewnorm(8.41, 3.04, 3.46) * 0.8;
  const double length = g.skewnorm(6.04, 3.02, 2.59) * 1;
  if (outflow) {
    R.blob0[0] = o.nblobs;
    base_jet(o, size, jet_amp, width, length, 4.0);
    R.nblob[0] = 1;
    R.strength[0] = g.normal(0.0, 0.2);
    R.radius[0] = g.uniform(50.0, 200.0);
  }
  double d0 = width / 2, d1 = length / 4;   // dims = np.sort([..]); dims[0] *= 1.2; dims = np.sort(dims)
  sort2(d0, d1);
  d0 *= 1.2;
  sort2(d0, d1);
  double c_sx = 0, c_sy = 0, c_rot = 0;
  bool vertical = true;
  for (int i = 0; i < ncomps; ++i) {
    const double dv0 = 1 + g.normal(0.0, 0.2), dv1 = 1 + g.normal(0.0, 0.2);
    double ox = 0, oy = 0;
    if (i != 0) { ox = g.uniform(-5.0, 5.0); oy = g.uniform(-5.0, 5.0); }
    double sx, sy;
    if (i == 0) {
      sx = bounded_width(g, d0, d1);
      sy = bounded_width(g, d0, d1);
      int tries = 0;
      while ((sx / sy < 0.33 || sx / sy > 3) || (0.75 < sx / sy && sx / sy < 1.33)) {
        if (tries > 10) {
          const double c0 = g.uniform(0.4, 0.8), c1 = g.uniform(1.2, 2.5);
          sy = sx * (g.integers(0, 2) == 0 ? c0 : c1);   // rng.choice([c0, c1])
          if (sy > 15) sy /= 5;
          break;
        }
        sy = bounded_width(g, d0, d1);
        ++tries;
      }
      vertical = sx > sy;
    } else {
      sx = c_sx * dv0;
      sy = c_sy * dv1;
    }
    double rot;
    const double pi = 3.141592653589793;
    if (i > 0) rot = g.uniform(c_rot - 0.1, c_rot + 0.1);
    else if (vertical) rot = g.uniform(-pi / 4, pi / 4);
    else rot = g.uniform(-pi / 4 + pi / 2, pi / 4 + pi / 2);
    const double c_amp = amp * g.uniform(0.5, 1.2);
    R.g[i] = Gauss{c_amp, size / 2.0 - ox, size / 2.0 - oy, sx, sy, rot};
    c_sx = sx; c_sy = sy; c_rot = rot;
  }
  R.angle = g.uniform(0.0, 360.0);
  R.compact = 1;
  R.amp = amp;
}

}  // namespace mjpack
}  // namespace rsb

using namespace rsb;
using namespace rsb::mjpack;

static int make_parent(const uint32_t* entropy, int n_entropy, const uint32_t* spawn_key, int n_spawn_key,
                       Parent* p) {
  RSB_ARG(entropy && n_entropy >= 1 && n_spawn_key >= 0 && (spawn_key || n_spawn_key == 0),
          "mojave pack: bad seed-sequence description");
  RSB_ARG(n_entropy + n_spawn_key + 4 <= kMaxKey, "mojave pack: entropy / spawn key too long");
  int n = 0;
  for (int i = 0; i < n_entropy; ++i) p->words[n++] = entropy[i];
  while (n < kPool) p->words[n++] = 0u;  // (a spawn key always follows: the child index)
  for (int i = 0; i < n_spawn_key; ++i) p->words[n++] = spawn_key[i];
  p->n = n;
  return 0;
}

extern "C" {

// Child generators of SeedSequence(entropy, spawn_key): PCG64 (state, inc) of children
// first_child .. first_child + n - 1 as numpy reports them: out[i] = {state hi, state lo, inc hi, inc lo}.
int rsb_pcg64_spawn_states(const uint32_t* entropy, int n_entropy, const uint32_t* spawn_key, int n_spawn_key,
                           long long first_child, int n, uint64_t* out) {
  Parent p;
  if (int e = make_parent(entropy, n_entropy, spawn_key, n_spawn_key, &p)) return e;
  RSB_ARG(out && n >= 0 && first_child >= 0 && first_child + n <= 0xFFFFFFFFLL, "rsb_pcg64_spawn_states: bad range");
  for (int i = 0; i < n; ++i) {
    const Pcg64 g = child_generator(p, (uint32_t)(first_child + i));
    out[4 * i] = (uint64_t)(g.state >> 64); out[4 * i + 1] = (uint64_t)g.state;
    out[4 * i + 2] = (uint64_t)(g.inc >> 64); out[4 * i + 3] = (uint64_t)g.inc;
  }
  return 0;
}

// `count` draws of one Generator method from PCG64 state {state hi, lo, inc hi, lo} (updated in place):
// kind 0 random(), 1 standard_normal(), 2 standard_exponential(), 3 power(a), 4 integers(lo=a, hi=b) as
// double, 5 uniform(a, b).  For the tests that compare this file's streams with numpy's.
int rsb_pcg64_draws(uint64_t* state4, int kind, double a, double b, int count, double* out) {
  RSB_ARG(state4 && out && count >= 0, "rsb_pcg64_draws: NULL argument");
  Pcg64 g;
  g.state = ((u128)state4[0] << 64) | state4[1];
  g.inc = ((u128)state4[2] << 64) | state4[3];
  g.has_uint32 = 0;
  g.uinteger = 0;
  for (int i = 0; i < count; ++i) {
    switch (kind) {
      case 0: out[i] = g.next_double(); break;
      case 1: out[i] = g.standard_normal(); break;
      case 2: out[i] = g.standard_exponential(); break;
      case 3: out[i] = g.power(a); break;
      case 4: out[i] = (double)g.integers((int64_t)a, (int64_t)b); break;
      case 5: out[i] = g.uniform(a, b); break;
      default: RSB_ARG(false, "rsb_pcg64_draws: unknown kind %d", kind);
    }
  }
  state4[0] = (uint64_t)(g.state >> 64); state4[1] = (uint64_t)g.state;
  return 0;
}

int rsb_mojave_pack(const uint32_t* entropy, int n_entropy, const uint32_t* spawn_key, int n_spawn_key,
                    long long first_child, int n, const int32_t* kinds, int size, void* recipes,
                    void* blob_params, int blob_cap, int32_t* n_blobs_out) {
  Parent p;
  if (int e = make_parent(entropy, n_entropy, spawn_key, n_spawn_key, &p)) return e;
  RSB_ARG(n >= 0 && kinds && recipes && blob_params && n_blobs_out && size >= 2 && blob_cap >= 0,
          "rsb_mojave_pack: bad arguments");
  RSB_ARG(first_child >= 0 && first_child + n <= 0xFFFFFFFFLL, "rsb_mojave_pack: child index out of range");
  Recipe* R = (Recipe*)recipes;
  Out o{(BlobParam*)blob_params, 0, blob_cap, false};
  for (int i = 0; i < n; ++i) {
    Pcg64 g = child_generator(p, (uint32_t)(first_child + i));
    std::memset(&R[i], 0, sizeof(Recipe));
    R[i].radius[0] = R[i].radius[1] = 100.0;
    switch (kinds[i]) {
      case 0: recipe_compact(g, size, R[i], o); break;
      case 1: recipe_one_jet(g, size, R[i], o); break;
      case 2: recipe_two_jet(g, size, R[i], o); break;
      default: RSB_ARG(false, "rsb_mojave_pack: kind %d of image %d (0 compact, 1 one jet, 2 two jets)", kinds[i], i);
    }
  }
  *n_blobs_out = o.nblobs;
  RSB_ARG(!o.overflow, "rsb_mojave_pack: %d blobs do not fit the buffer of %d", o.nblobs, blob_cap);
  return 0;
}

}  // extern "C"
