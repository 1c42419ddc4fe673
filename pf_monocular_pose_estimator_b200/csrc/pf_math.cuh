// pf_math.cuh -- per-particle math of the PF hot path (FP32 state, exact integer RNG).
//
// Everything here is `PF_HD` (host+device): the CUDA kernels in pf_kernels.cuh inline these
// functions; tests/ additionally compile this header with g++ (tests/hostmath.cpp) to check
// the device arithmetic against the oracle without a GPU.  That host build is a test
// artefact only -- the shipped library has no CPU execution path.
//
// Reference semantics restated (file:line are /root/reference/pf_mpe_lib/src/pose_estimator.cpp):
//   motion model            :543-588      pf_propagate
//   project2d               :1017-1034    pf_project
//   calculateEstimationProbability :2385-2445   pf_score
//   RNG stream              :476-477,:510-523 (libstdc++ minstd_rand0 + generate_canonical)
//   stratified resampling   :668-682      pf_threshold / pf_count_leq (source-centric form)
#pragma once

#include <stdint.h>
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define PF_HD __host__ __device__ __forceinline__
#else
#define PF_HD inline
#endif
#if defined(__CUDACC__)
#define PF_UNROLL _Pragma("unroll")
#else
#define PF_UNROLL
#endif

#define PF_MAX_LED 8
#define PF_RNG_M 2147483647u     /* 2^31 - 1 */
#define PF_RNG_A 16807u
#define PF_RNG_ORDER 2147483646u /* multiplicative order divides m-1 */
#define PF_FX_BITS 30            /* weights are accumulated as Q33.30 fixed point (int64) */

struct pf_f2 { float x, y; };

// ----------------------------------------------------------------------------------------
// RNG: x <- 16807 x mod (2^31-1); exact skip-ahead with radix-256 power tables.
// ----------------------------------------------------------------------------------------
PF_HD uint32_t pf_modmul(uint32_t a, uint32_t b)
{
  uint64_t p = (uint64_t)a * (uint64_t)b;                 // < 2^62
  uint32_t s = (uint32_t)(p & PF_RNG_M) + (uint32_t)(p >> 31);
  s = (s & PF_RNG_M) + (s >> 31);
  return s >= PF_RNG_M ? s - PF_RNG_M : s;
}

// one engine step: x * 16807 < 2^46, so a single fold leaves s < 2^31 + 2^15 and one conditional
// subtraction (as an unsigned min) finishes the reduction
PF_HD uint32_t pf_rng_next(uint32_t x)
{
  const uint64_t p = (uint64_t)x * (uint64_t)PF_RNG_A;
  const uint32_t s = (uint32_t)(p & PF_RNG_M) + (uint32_t)(p >> 31);
  const uint32_t t = s - PF_RNG_M;          // wraps to a huge value when s < m
  return s < t ? s : t;
}

// two engine steps at once: x * 16807^2 mod m.  16807^2 = 282475249 < 2^29, so the product is < 2^60, one fold leaves
// s < 2^31 + 2^29 and one conditional subtraction finishes (s == m would need x == 0 mod m: impossible for a live state)
#define PF_RNG_A2 282475249u
PF_HD uint32_t pf_rng_next2(uint32_t x)
{
  const uint64_t p = (uint64_t)x * (uint64_t)PF_RNG_A2;
  const uint32_t s = (uint32_t)(p & PF_RNG_M) + (uint32_t)(p >> 31);
  const uint32_t t = s - PF_RNG_M;
  return s < t ? s : t;
}

// Skip-ahead: radix-512 power tables, tab[d*512 + v] = 16807^(v * 512^d) mod m, d = 0..3 (36 exponent bits >= 31);
// x * 16807^e = one table entry + one modular multiplication per 9-bit digit of e.  Exponents are taken mod
// PF_RNG_ORDER (16807 is a primitive root of 2^31-1).
#define PF_RNG_DIGIT_BITS 9
#define PF_RNG_TAB_WORDS (4 << PF_RNG_DIGIT_BITS)
inline void pf_build_rng_table(uint32_t* tab)     // host
{
  const int R = 1 << PF_RNG_DIGIT_BITS;
  uint32_t base = PF_RNG_A;
  for (int d = 0; d < 4; ++d) {
    tab[d * R] = 1u;
    for (int v = 1; v < R; ++v) tab[d * R + v] = pf_modmul(tab[d * R + v - 1], base);
    base = pf_modmul(tab[d * R + R - 1], base);   // base^R
  }
}
// nd = number of digits that can be non-zero (uniform per call site: the exponent is below 512^nd)
PF_HD uint32_t pf_rng_skip_nd(uint32_t x, uint32_t e, const uint32_t* tab, int nd)
{
  const uint32_t M9 = (1u << PF_RNG_DIGIT_BITS) - 1u;
  x = pf_modmul(x, tab[e & M9]);
  x = pf_modmul(x, tab[(1 << PF_RNG_DIGIT_BITS) + ((e >> PF_RNG_DIGIT_BITS) & M9)]);
  if (nd > 2) x = pf_modmul(x, tab[(2 << PF_RNG_DIGIT_BITS) + ((e >> (2 * PF_RNG_DIGIT_BITS)) & M9)]);
  if (nd > 3) x = pf_modmul(x, tab[(3 << PF_RNG_DIGIT_BITS) + (e >> (3 * PF_RNG_DIGIT_BITS))]);
  return x;
}
// pf_rng_skip_e: exponent already reduced mod PF_RNG_ORDER
PF_HD uint32_t pf_rng_skip_e(uint32_t x, uint32_t e, const uint32_t* tab) { return pf_rng_skip_nd(x, e, tab, 4); }
#define PF_RNG_A384 1618907920u /* 16807^384 mod (2^31-1): 32 particles x 12 engine steps (checked in pfmpe_api.cu) */

PF_HD uint32_t pf_rng_skip(uint32_t x, uint64_t steps, const uint32_t* __restrict__ tab)
{
  return pf_rng_skip_nd(x, (uint32_t)(steps % PF_RNG_ORDER), tab, 4);
}

// generate_canonical<double,53> from two consecutive engine outputs, bit-exact with libstdc++
// (sum = (g1-1) + (g2-1)*R; ret = sum / (R*R)), no FMA contraction.
PF_HD double pf_canonical_f64(uint32_t g1, uint32_t g2)
{
  const double R = 2147483646.0;
  const double RR = 2147483646.0 * 2147483646.0;  // constant-folded with one IEEE rounding
#if defined(__CUDA_ARCH__)
  double sum = __dadd_rn((double)(g1 - 1u), __dmul_rn((double)(g2 - 1u), R));
  double c = __ddiv_rn(sum, RR);
#else
  volatile double prod = (double)(g2 - 1u) * R;   // volatile: forbid host FMA contraction
  double sum = (double)(g1 - 1u) + prod;
  double c = sum / RR;
#endif
  if (c >= 1.0) c = 0.99999999999999988897769753748;  // nextafter(1,0)
  return c;
}

// FP32 rendering of the same canonical value (within 1 ulp_f32 of the double one): used for
// the motion noise, whose consumer is FP32 anyway.
PF_HD float pf_canonical_f32(uint32_t g1, uint32_t g2)
{
  const float invR = 1.0f / 2147483646.0f;
  float hi = (float)(g2 - 1u), lo = (float)(g1 - 1u);
  return fminf(fmaf(lo, invR, hi) * invR, 0.99999994f);
}

// The motion noise only needs the high engine output of a draw: generate_canonical's value is
// ((g1-1) + (g2-1) R) / R^2 = (g2-1)/R + (g1-1)/R^2, and the second term is below 2^-31 -- under half an ulp_f32 of
// every value the first term can take except the few below 2^-7, where it still moves the noise by less than 1e-11 of
// its range.  The engine is stepped twice per draw all the same (pf_rng_next2), so the stream stays the reference's.
PF_HD float pf_canonical_f32_hi(uint32_t g2)
{
  const float invR = 1.0f / 2147483646.0f;
  return fminf((float)(g2 - 1u) * invR, 0.99999994f);
}

// Fast reciprocal / square root of the scoring path.  Device: the SFU approximation (1 ulp); the reciprocal gets one
// Newton step (it scales pixel coordinates of several hundred, where an ulp is 3e-5 px: the 1e-5 bar on the weights
// leaves no room for a second ulp), the square root does not (it yields distances below the gate, a few pixels, whose
// ulp is 5e-7 px).  No slow-path branches: a zero / denormal / infinite argument gives inf, 0 or NaN, which the callers
// treat like the IEEE results (a NaN or inf distance is never within the gate).
PF_HD float pf_rcp_nr(float x)
{
#if defined(__CUDA_ARCH__)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return fmaf(fmaf(-x, r, 1.0f), r, r);
#else
  return 1.0f / x;
#endif
}
PF_HD float pf_sqrt_fast(float x)
{
#if defined(__CUDA_ARCH__)
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return sqrtf(x);
#endif
}

// ----------------------------------------------------------------------------------------
// Poses: 3x4 row-major float, implicit bottom row (0,0,0,1).  p[4*i+j]; p[4*i+3] = t_i.
// ----------------------------------------------------------------------------------------
struct pf_pose { float p[12]; };

PF_HD pf_pose pf_affine_mul(const pf_pose& A, const pf_pose& B)
{
  pf_pose C;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float s = A.p[4 * i] * B.p[j];
      s = fmaf(A.p[4 * i + 1], B.p[4 + j], s);
      s = fmaf(A.p[4 * i + 2], B.p[8 + j], s);
      if (j == 3) s += A.p[4 * i + 3];
      C.p[4 * i + j] = s;
    }
  }
  return C;
}

struct pf_noise_cfg {
  float ang_lo, ang_span;  // U(a, b) = c*(b-a)+a  with a = minAngularNoise*facRot   (:514-516)
  float tr_lo, tr_span;    // same for the translation ranges                          (:510-512)
};

// Motion model for particle n >= 2 (:555-587).  `base` = camMoveInv*E_n[*predictionMatrix];
// x = engine state right before this particle's first draw; 12 engine steps are consumed in
// the order aX,aY,aZ,tX,tY,tZ.  scale = 1 + 0.025*floor(iter/10).
// sin/cos for |x| <= 0.3 (truncation error < 1e-12 relative: below FP32 rounding)
PF_HD void pf_sincos_small(float x, float& s, float& c)
{
  const float x2 = x * x;
  float p = fmaf(x2, 2.7557319e-6f, -1.9841270e-4f);
  p = fmaf(p, x2, 8.3333333e-3f);
  p = fmaf(p, x2, -1.6666667e-1f);
  s = fmaf(x * x2, p, x);
  float q = fmaf(x2, 2.4801587e-5f, -1.3888889e-3f);
  q = fmaf(q, x2, 4.1666667e-2f);
  q = fmaf(q, x2, -0.5f);
  c = fmaf(q, x2, 1.0f);
}

// sin/cos for |x| <= 0.03 (truncation: x^5/120 < 2.1e-10, x^6/720 < 1.1e-12 absolute)
PF_HD void pf_sincos_tiny(float x, float& s, float& c)
{
  const float x2 = x * x;
  s = x * fmaf(x2, -1.6666667e-1f, 1.0f);
  c = fmaf(fmaf(x2, 4.1666667e-2f, -0.5f), x2, 1.0f);
}

// ang_mode: 0 = libm sincosf, 1 = every possible draw within +-0.3 rad (PF_FC_SMALL_ANG), 2 = within +-0.03 rad
// (PF_FC_TINY_ANG: the default noise ranges, +-0.015 rad x facRot x scale).
// The six noise values are  (c_d * span + lo) * scale  with c_d = (g2 - 1) / R (pf_canonical_f32_hi); with the two
// uniform factors folded that is one conversion and one FMA per draw: g2 * (span * scale / R) + lo * scale.  (g2 - 1
// and the clamp below 1 of the canonical value are dropped with the same argument as g1: they move the noise by
// less than 2^-31 of its range.)
PF_HD pf_pose pf_propagate(const pf_pose& base, uint32_t x, const pf_noise_cfg& nc, float scale, int ang_mode)
{
  const float invR = 1.0f / 2147483646.0f;
  const float ka = nc.ang_span * scale * invR, la = nc.ang_lo * scale;
  const float kt = nc.tr_span * scale * invR, lt = nc.tr_lo * scale;
  float c6[6];
#pragma unroll
  for (int d = 0; d < 6; ++d) {
    x = pf_rng_next2(x);             // g2 of this draw (g1 = the step in between is not needed, see pf_canonical_f32_hi)
    c6[d] = (float)x;
  }
  const float a = fmaf(c6[0], ka, la);
  const float b = fmaf(c6[1], ka, la);
  const float c = fmaf(c6[2], ka, la);
  float sa, ca, sb, cb, sc, cc;
  if (ang_mode == 2) {
    pf_sincos_tiny(a, sa, ca); pf_sincos_tiny(b, sb, cb); pf_sincos_tiny(c, sc, cc);
  } else if (ang_mode == 1) {
    pf_sincos_small(a, sa, ca); pf_sincos_small(b, sb, cb); pf_sincos_small(c, sc, cc);
  } else {
#if defined(__CUDA_ARCH__)
    sincosf(a, &sa, &ca); sincosf(b, &sb, &cb); sincosf(c, &sc, &cc);
#else
    sa = sinf(a); ca = cosf(a); sb = sinf(b); cb = cosf(b); sc = sinf(c); cc = cosf(c);
#endif
  }
  // Rn = Rz(c) * Ry(b) * Rx(a)   (:567-582)
  float r[9];
  r[0] = cc * cb; r[1] = cc * sb * sa - sc * ca; r[2] = cc * sb * ca + sc * sa;
  r[3] = sc * cb; r[4] = sc * sb * sa + cc * ca; r[5] = sc * sb * ca - cc * sa;
  r[6] = -sb;     r[7] = cb * sa;                r[8] = cb * ca;
  pf_pose T;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      float s = base.p[4 * i] * r[j];
      s = fmaf(base.p[4 * i + 1], r[3 + j], s);
      s = fmaf(base.p[4 * i + 2], r[6 + j], s);
      T.p[4 * i + j] = s;
    }
  }
  // translation added separately (:585-587)
  T.p[3]  = base.p[3]  + fmaf(c6[3], kt, lt);
  T.p[7]  = base.p[7]  + fmaf(c6[4], kt, lt);
  T.p[11] = base.p[11] + fmaf(c6[5], kt, lt);
  return T;
}

// project2d (:1017-1034): p = K (R X + t), (u,v) = (p0/p2, p1/p2); full 3x3 K, no cheirality test.
// kstd (uniform): K has the usual zero pattern (K01 = K10 = K20 = K21 = 0, K22 = 1); the skipped
// terms are exact zeros, so both branches give the same bits for finite inputs.  The two
// divisions share one reciprocal (pf_rcp_nr: <= 1 ulp from the quotient).
PF_HD bool pf_k_is_std(const float* K9) { return K9[1] == 0.0f && K9[3] == 0.0f && K9[6] == 0.0f && K9[7] == 0.0f && K9[8] == 1.0f; }

PF_HD void pf_to_camera(const pf_pose& T, float X, float Y, float Z, float& xc, float& yc, float& zc)
{
  xc = fmaf(T.p[0], X, fmaf(T.p[1], Y, fmaf(T.p[2], Z, T.p[3])));
  yc = fmaf(T.p[4], X, fmaf(T.p[5], Y, fmaf(T.p[6], Z, T.p[7])));
  zc = fmaf(T.p[8], X, fmaf(T.p[9], Y, fmaf(T.p[10], Z, T.p[11])));
}
PF_HD void pf_to_pixel(const float* __restrict__ K9, bool kstd, float xc, float yc, float zc, float& u, float& v)
{
  float p0, p1, p2;
  if (kstd) {
    p0 = fmaf(K9[0], xc, K9[2] * zc);
    p1 = fmaf(K9[4], yc, K9[5] * zc);
    p2 = zc;
  } else {
    p0 = fmaf(K9[0], xc, fmaf(K9[1], yc, K9[2] * zc));
    p1 = fmaf(K9[3], xc, fmaf(K9[4], yc, K9[5] * zc));
    p2 = fmaf(K9[6], xc, fmaf(K9[7], yc, K9[8] * zc));
  }
  const float r = pf_rcp_nr(p2);
  u = p0 * r;
  v = p1 * r;
}
PF_HD void pf_project(const float* __restrict__ K9, bool kstd, const pf_pose& T, float X, float Y, float Z, float& u, float& v)
{
  float xc, yc, zc;
  pf_to_camera(T, X, Y, Z, xc, yc, zc);
  pf_to_pixel(K9, kstd, xc, yc, zc, u, v);
}

// Forward plumb_bob distortion of a projected pixel (row f3): the model of LEDDetector::distortPoints
// (led_detector.cpp:371-414) -- normalise with K, radial (k1, k2, k3) and tangential (p1, p2) terms, back
// to pixels -- so that particles can be scored against the raw, still-distorted blob centres.
// d = (k1, k2, p1, p2, k3), the CameraInfo.D order.
PF_HD void pf_distort(const float* K9, const float* d, float& u, float& v)
{
  const float fx = K9[0], fy = K9[4], cx = K9[2], cy = K9[5];
  const float x = (u - cx) / fx, y = (v - cy) / fy;
  const float r2 = x * x + y * y;
  const float rad = 1.0f + d[0] * r2 + d[1] * r2 * r2 + d[4] * r2 * r2 * r2;
  const float xc = x * rad + (2.0f * d[2] * x * y + d[3] * (r2 + 2.0f * x * x));
  const float yc = y * rad + (d[2] * (r2 + 2.0f * y * y) + 2.0f * d[3] * x * y);
  u = xc * fx + cx;
  v = yc * fy + cy;
}

// ----------------------------------------------------------------------------------------
// Scoring (:2385-2445) in the per-LED-minimum form: for every LED the nearest
// blob (ties -> smallest blob index), then LEDs in ascending (min, j) order, at most
// min(D,L), stop at the first beyond tol_PF.
// ----------------------------------------------------------------------------------------
struct pf_score_cfg {
  float tol, inv_tol;   // back_projection_pixel_tolerance_ (gain)        :2416
  float tol_pf;         // back_projection_pixel_tolerance_PF (gate)      :2414
  unsigned dg_mask;     // bMarkerDowngrade bits (first five LEDs only)   :2431
  float led_base[PF_MAX_LED];   // L, or L - 2 for a downgraded LED: the constant part of an accepted LED's term (:2416, :2431-2432)
};
PF_HD void pf_score_cfg_finish(pf_score_cfg& sc, int L)
{
  for (int j = 0; j < PF_MAX_LED; ++j) sc.led_base[j] = (float)((j < 5 && ((sc.dg_mask >> j) & 1u)) ? L - 2 : L);
}

// Greedy part (:2409-2438) on the per-LED nearest-blob results (m[j] = squared distance >= 0 or
// +inf, am[j] = blob index < 65536).  The reference takes the LEDs in ascending (distance, LED)
// order: the keys (bits(m) << 32 | j << 16 | am) are sorted with a fixed compare-exchange network
// (squared distances are non-negative, so their bit patterns order like the values) and walked
// until the first one beyond the gate.
template <int L, bool WANT_PAIRS>
PF_HD float pf_greedy(const float* m, const int* am, int D, const pf_score_cfg& sc, unsigned* pairs, int* npairs)
{
  uint64_t key[L];
#pragma unroll
  for (int j = 0; j < L; ++j) {
    uint32_t mb;
#if defined(__CUDA_ARCH__)
    mb = __float_as_uint(m[j]);
#else
    { float t = m[j]; memcpy(&mb, &t, 4); }
#endif
    key[j] = ((uint64_t)mb << 32) | ((uint64_t)j << 16) | (uint64_t)(unsigned)am[j];
  }
  // odd-even transposition sort, L rounds (all indices compile-time)
#pragma unroll
  for (int r = 0; r < L; ++r) {
#pragma unroll
    for (int i = (r & 1); i + 1 < L; i += 2) {
      const uint64_t a = key[i], b = key[i + 1];
      key[i] = a < b ? a : b;
      key[i + 1] = a < b ? b : a;
    }
  }
  const int steps = D < L ? D : L;   // :2409
  float P = 0.0f;
  int occ = 1, nacc = 0;
  unsigned prev_am[L];
#pragma unroll
  for (int k = 0; k < L; ++k) {
    if (k >= steps) break;
    const uint32_t mb = (uint32_t)(key[k] >> 32);
    const unsigned best = (unsigned)(key[k] >> 16) & 0xFFFFu, bam = (unsigned)key[k] & 0xFFFFu;
    float bm;
#if defined(__CUDA_ARCH__)
    bm = __uint_as_float(mb);
#else
    memcpy(&bm, &mb, 4);
#endif
    const float d = sqrtf(bm);       // :2412
    if (!(d <= sc.tol_pf)) break;    // :2414 / :2438
    const float g = (sc.tol - d) * sc.inv_tol;
    P += (float)L + g * g;           // :2416
    bool reused = false;             // :2421
#pragma unroll
    for (int q = 0; q < k; ++q) reused = reused || (prev_am[q] == bam);
    prev_am[k] = bam;
    if (reused) { P -= (float)(3 * occ); ++occ; }   // :2423-2424
    if ((sc.dg_mask >> best) & 1u) P -= 2.0f;       // :2431-2432
    if (WANT_PAIRS) { pairs[2 * nacc] = best + 1u; pairs[2 * nacc + 1] = bam + 1u; }
    ++nacc;
  }
  if (WANT_PAIRS) *npairs = nacc;
  return P;
}

// The weight alone, without the sort.  The greedy loop accepts LEDs in ascending distance order until the first one
// beyond the gate: the accepted set is simply {j : d_j <= tol_PF}, every accepted LED contributes L + g_j^2 (- 2 if
// downgraded), and the shared-blob penalty 3 + 6 + ... (:2421-2424) depends only on HOW MANY accepted LEDs find their
// blob already taken by an earlier one = accepted - distinct blobs among them, whatever the order.  What the order does
// change is the FP32 rounding of the sum (terms are added in LED order here), far below the 1e-5 bar against the FP64
// reference.  Used by the score pass of every particle; the winner's correspondences still come from pf_greedy.
template <int L>
PF_HD float pf_weight_unordered(const float* m, const int* am, int steps, const pf_score_cfg& sc)
{
  // steps = min(D, L) (:2409).  With fewer detections than LEDs only the `steps` nearest LEDs (ascending distance, ties
  // to the smaller LED index) are looked at: rank_j = number of LEDs that precede j in that order.
  int rank[L];
  PF_UNROLL
  for (int j = 0; j < L; ++j) rank[j] = 0;
  if (steps < L) {
    PF_UNROLL
    for (int i = 0; i < L; ++i) {
      PF_UNROLL
      for (int j = i + 1; j < L; ++j) {
        const bool i_first = m[i] <= m[j] || !(m[j] == m[j]);   // NaN distances sort last
        rank[j] += i_first ? 1 : 0;
        rank[i] += i_first ? 0 : 1;
      }
    }
  }
  float P = 0.0f;
  int key[L];
  PF_UNROLL
  for (int j = 0; j < L; ++j) {
    const float d = pf_sqrt_fast(m[j]);                                  // :2412
    const bool acc = d <= sc.tol_pf && rank[j] < steps;                  // :2414 (NaN / inf: not accepted)
    const float g = (sc.tol - d) * sc.inv_tol;
    const float t = fmaf(g, g, sc.led_base[j]);                          // :2416, :2431-2432
    P += acc ? t : 0.0f;
    key[j] = acc ? am[j] : -1 - j;                                       // rejected LEDs never match anything
  }
  int dup = 0;
  PF_UNROLL
  for (int k = 1; k < L; ++k) {
    bool taken = false;
    PF_UNROLL
    for (int q = 0; q < k; ++q) taken = taken || (key[q] == key[k]);
    dup += taken ? 1 : 0;
  }
  return P - 1.5f * (float)(dup * (dup + 1));                            // 3 * (1 + 2 + ... + dup), :2423-2424
}

template <int L, bool WANT_PAIRS>
PF_HD float pf_score(const pf_f2* __restrict__ blobs, int D, const float* u, const float* v,
                     const pf_score_cfg& sc, unsigned* pairs, int* npairs)
{
  float m[L];
  int am[L];
#pragma unroll
  for (int j = 0; j < L; ++j) { m[j] = INFINITY; am[j] = 0; }
  for (int i = 0; i < D; ++i) {
    const pf_f2 z = blobs[i];
#pragma unroll
    for (int j = 0; j < L; ++j) {
      float dx = z.x - u[j], dy = z.y - v[j];
      float d2 = fmaf(dy, dy, dx * dx);
      if (d2 < m[j]) { m[j] = d2; am[j] = i; }
    }
  }
  if (!WANT_PAIRS) return pf_weight_unordered<L>(m, am, D < L ? D : L, sc);
  return pf_greedy<L, WANT_PAIRS>(m, am, D, sc, pairs, npairs);
}

// ----------------------------------------------------------------------------------------
// Fixed-point weights and stratified-resampling thresholds.
//
// Raw FP32 scores are accumulated exactly as int64 Q.30, so sums / prefix sums are associative:
// any tiling, any number of CTAs or GPUs gives bit-identical totals and prefixes.  With
// S = sum_i w_i (raw), sg = sign(S), the reference test  (sum_{j<=i} w_j / S) >= u_k  (:676-677)
// becomes  sg * prefix_fx(i) >= t_k,  t_k = ceil(u_k * |S_fx|).
// ----------------------------------------------------------------------------------------
PF_HD long long pf_w_to_fx(float w)
{
#if defined(__CUDA_ARCH__)
  return __float2ll_rn(w * 1073741824.0f);     // the scaling is exact (power of two), one rounding: same value as the FP64 form
#else
  return (long long)llrint((double)w * 1073741824.0);
#endif
}

struct pf_resample_cfg {
  uint32_t x_base;   // engine state right before resample draw 0 (after 12*(N-2)*I steps)
  int N;             // number of strata (global N when sharded)
  double Nd;         // (double)N
  double Sabs;       // (double)|S_fx|
  double NdS;        // Nd / Sabs (estimate only; the threshold test itself is exact)
  double margin;     // fast-path guard, see pf_count_leq
  double invN;       // 1 / Nd (estimate only)
  int nd;            // radix-512 digits of the largest skip exponent 2 (N - 1) (pf_rng_skip_nd)
};

PF_HD pf_resample_cfg pf_make_resample_cfg(uint32_t x_base, int N, long long S_fx)
{
  pf_resample_cfg c;
  c.x_base = x_base;
  c.N = N;
  c.Nd = (double)N;
  long long a = S_fx < 0 ? -S_fx : S_fx;
  c.Sabs = (double)a;
  c.NdS = c.Nd / (c.Sabs > 0 ? c.Sabs : 1.0);
  c.margin = 8.0 * 2.220446049250313e-16 * c.Nd + 2.0 * c.NdS + 1e-12;
  c.invN = 1.0 / c.Nd;
  c.nd = N <= (1 << 17) ? 2 : (N <= (1 << 26) ? 3 : 4);
  return c;
}

// t_k = ceil( ((k + r_k)/N) * |S_fx| ), r_k = draw k of randResample (:671), double ops as in
// the reference ((k + r)/N is an IEEE add and an IEEE divide).  Non-decreasing in k.
#if defined(__CUDA_ARCH__)
#define PF_HD_COLD __device__ __noinline__     /* rare paths: one out-of-line copy keeps the hot loop's code small */
#else
#define PF_HD_COLD inline
#endif
PF_HD_COLD long long pf_threshold(int k, const pf_resample_cfg& rc, const uint32_t* __restrict__ tab)
{
  uint32_t x = pf_rng_skip(rc.x_base, 2ull * (uint64_t)k, tab);
  uint32_t g1 = pf_rng_next(x), g2 = pf_rng_next(g1);
  double r = pf_canonical_f64(g1, g2);
#if defined(__CUDA_ARCH__)
  double u = __ddiv_rn(__dadd_rn((double)k, r), rc.Nd);
  double td = __dmul_rn(u, rc.Sabs);
#else
  volatile double s = (double)k + r;
  volatile double u = s / rc.Nd;
  double td = u * rc.Sabs;
#endif
  return (long long)ceil(td);
}

// K(m) = #{k in [0,N): t_k <= m}.  Offspring of source i are the strata [K(M_{i-1}), K(M_i))
// where M is the running maximum of sg*prefix_fx (first-crossing semantics for non-monotone
// prefixes).  One RNG evaluation in the common case.
PF_HD int pf_count_leq(long long m, const pf_resample_cfg& rc, const uint32_t* __restrict__ tab)
{
  if (m < 0) return 0;
  double x = (double)m * rc.NdS;
  int kf;
  if (!(x < rc.Nd)) kf = rc.N; else kf = (int)x;   // floor, x >= 0
  double frac = x - (double)kf;
  if (kf < rc.N && frac > rc.margin && frac < 1.0 - rc.margin) {
    // The answer is kf + [t_kf <= m].  t_kf = ceil(td) with td = ((kf + r) / N) * |S|, and for the integer m
    // ceil(td) <= m  <=>  td <= m.  td is first estimated with reciprocal multiplications instead of the two
    // divisions (relative error below 1e-15: six roundings + two rounded constants); the estimate decides unless it
    // lies within 4e-15 relative of m, in which case the exact, division-based threshold does.
    const uint32_t xs = pf_rng_skip_nd(rc.x_base, 2u * (uint32_t)kf, tab, rc.nd);   // 2 kf < 2^31 - 2: no reduction needed
    const uint32_t g1 = pf_rng_next(xs), g2 = pf_rng_next(g1);
    const double sum = (double)(g1 - 1u) + (double)(g2 - 1u) * 2147483646.0;
    const double r = sum * (1.0 / (2147483646.0 * 2147483646.0));
    const double td = ((double)kf + r) * rc.invN * rc.Sabs;
    const double md = (double)m, band = td * 4.0e-15;
    if (td < md - band) return kf + 1;
    if (td > md + band) return kf;
    return kf + (pf_threshold(kf, rc, tab) <= m ? 1 : 0);
  }
  int K = kf;
  while (K < rc.N && pf_threshold(K, rc, tab) <= m) ++K;
  while (K > 0 && pf_threshold(K - 1, rc, tab) > m) --K;
  return K;
}

// (sum, running-max-of-prefix) scan operator:  (s1,m1) (+) (s2,m2) = (s1+s2, max(m1, s1+m2)).
struct pf_sm { long long s, m; };
#define PF_SM_NEG (-(1ll << 62))
#define PF_SM_POS (1ll << 62)

// Sign-agnostic aggregate of a run of particles, built while scoring (before the sign of the
// weight sum is known): sum, running max and running min of the prefix.  After the
// grid-wide sum S is known, sg = sign(S) selects  M = sg > 0 ? mx : -mn.
struct pf_sm3 { long long s, mx, mn; };
PF_HD pf_sm3 pf_sm3_identity() { pf_sm3 r; r.s = 0; r.mx = PF_SM_NEG; r.mn = PF_SM_POS; return r; }
PF_HD pf_sm3 pf_sm3_combine(const pf_sm3& a, const pf_sm3& b)
{
  pf_sm3 r;
  r.s = a.s + b.s;
  const long long tx = b.mx <= PF_SM_NEG ? PF_SM_NEG : a.s + b.mx;
  const long long tn = b.mn >= PF_SM_POS ? PF_SM_POS : a.s + b.mn;
  r.mx = a.mx > tx ? a.mx : tx;
  r.mn = a.mn < tn ? a.mn : tn;
  return r;
}
// (sum, running max) of sg * prefix from the sign-agnostic aggregate
PF_HD void pf_sm3_select(const pf_sm3& a, long long sg, long long& s, long long& m)
{
  s = sg * a.s;
  if (sg > 0) m = a.mx <= PF_SM_NEG ? PF_SM_NEG : a.mx;
  else m = a.mn >= PF_SM_POS ? PF_SM_NEG : -a.mn;
}
PF_HD pf_sm pf_sm_identity() { pf_sm r; r.s = 0; r.m = PF_SM_NEG; return r; }
PF_HD pf_sm pf_sm_combine(const pf_sm& a, const pf_sm& b)
{
  pf_sm r;
  r.s = a.s + b.s;
  long long t = b.m <= PF_SM_NEG ? PF_SM_NEG : a.s + b.m;
  r.m = a.m > t ? a.m : t;
  return r;
}
