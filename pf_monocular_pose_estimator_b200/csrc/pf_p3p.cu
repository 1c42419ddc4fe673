// pf_p3p.cu -- kernels of K9; see pf_p3p.cuh.  Built with -fmad=false.
#include "pf_p3p.cuh"

#include <cstdlib>

struct pf_cxd { double re, im; };

__device__ __forceinline__ pf_cxd pf_cx(double re, double im) { pf_cxd z; z.re = re; z.im = im; return z; }
__device__ __forceinline__ pf_cxd pf_cx_add(pf_cxd a, pf_cxd b) { return pf_cx(a.re + b.re, a.im + b.im); }
__device__ __forceinline__ pf_cxd pf_cx_sub(pf_cxd a, pf_cxd b) { return pf_cx(a.re - b.re, a.im - b.im); }
__device__ __forceinline__ pf_cxd pf_cx_neg(pf_cxd a) { return pf_cx(-a.re, -a.im); }
__device__ __forceinline__ pf_cxd pf_cx_scale(pf_cxd a, double s) { return pf_cx(a.re * s, a.im * s); }
__device__ __forceinline__ pf_cxd pf_cx_divs(pf_cxd a, double s) { return pf_cx(a.re / s, a.im / s); }

// complex quotient (Smith's scaling, as the compiler runtime's complex division does for finite operands)
__device__ __noinline__ pf_cxd pf_cx_div(pf_cxd a, pf_cxd b)
{
  double re, im;
  if (fabs(b.re) < fabs(b.im)) {
    const double r = b.re / b.im, den = b.re * r + b.im;
    re = (a.re * r + a.im) / den;
    im = (a.im * r - a.re) / den;
  } else {
    const double r = b.im / b.re, den = b.im * r + b.re;
    re = (a.im * r + a.re) / den;
    im = (a.im - a.re * r) / den;
  }
  return pf_cx(re, im);
}

// principal square root with the branch rules of C99 csqrt (the sign of a zero imaginary part selects the branch)
__device__ __noinline__ pf_cxd pf_cx_sqrt(pf_cxd z)
{
  if (z.re != z.re || z.im != z.im) return pf_cx(z.re + z.im, z.re + z.im);
  if (z.im == 0.0) {
    if (z.re < 0.0) return pf_cx(0.0, copysign(sqrt(-z.re), z.im));
    return pf_cx(fabs(sqrt(z.re)), copysign(0.0, z.im));
  }
  if (z.re == 0.0) {
    const double r = sqrt(0.5 * fabs(z.im));
    return pf_cx(r, copysign(r, z.im));
  }
  const double d = hypot(z.re, z.im);
  double r, s;
  if (z.re > 0.0) {
    r = sqrt(0.5 * (d + z.re));
    s = 0.5 * (z.im / r);
  } else {
    s = sqrt(0.5 * (d - z.re));
    r = fabs(0.5 * (z.im / s));
  }
  return pf_cx(r, copysign(s, z.im));
}

// std::pow(std::complex<double>, double) as libstdc++ defines it: a positive real base stays real, everything
// else goes through polar(exp(y * log|z|), y * arg z) -- which gives a negative real base a tiny imaginary part
// cube_root: y is fl(1/3); for a positive real base cbrt() stands in for pow(x, fl(1/3)) -- the two agree to
// within an ulp (x^(fl(1/3) - 1/3) = 1 - 2e-17 ln x) and cbrt is a fraction of pow's code and time.
__device__ __noinline__ pf_cxd pf_cx_pow(pf_cxd z, double y, bool cube_root)
{
  if (z.im == 0.0 && z.re > 0.0) return pf_cx(cube_root ? cbrt(z.re) : pow(z.re, y), 0.0);
  const double lr = log(hypot(z.re, z.im)), th = atan2(z.im, z.re);
  const double rho = exp(y * lr);
  double sn, cs;
  sincos(y * th, &sn, &cs);
  return pf_cx(rho * cs, rho * sn);
}

// pow(complex(x, +0), n) for n = 2, 3 under the same rules: positive x stays real (x^n is formed by multiplication:
// within an ulp of pow); otherwise arg = pi exactly, so the polar form is exp(n log|x|) * (cos, sin)(n * fl(pi)),
// two constants: cos(2 fl(pi)) = 1, sin(2 fl(pi)) = -2.4492935982947064e-16, cos(3 fl(pi)) = -1,
// sin(3 fl(pi)) = 3.6739403974420594e-16 (this is where the reference's "real" quartic gets its imaginary dust).
__device__ __forceinline__ pf_cxd pf_cx_pow_real_int(double x, int n)
{
  if (x > 0.0) return pf_cx(n == 2 ? x * x : x * x * x, 0.0);
  if (x == 0.0 && !signbit(x)) return pf_cx(0.0, 0.0);   // arg(+0 + 0i) = 0
  const double rho = exp((double)n * log(fabs(x)));
  return n == 2 ? pf_cx(rho * 1.0, rho * -2.4492935982947064e-16) : pf_cx(rho * -1.0, rho * 3.6739403974420594e-16);
}

// P3P::solveQuartic (p3p.cpp:238-292): Ferrari's closed form, real parts of the four roots
__device__ __forceinline__ void pf_p3p_quartic(const double f[5], double roots[4])
{
  const double A = f[0], B = f[1], C = f[2], D = f[3], E = f[4];
  const double A2 = A * A, B2 = B * B, A3 = A2 * A, B3 = B2 * B, A4 = A3 * A, B4 = B3 * B;
  const double alpha = -3 * B2 / (8 * A2) + C / A;
  const double beta = B3 / (8 * A3) - B * C / (2 * A2) + D / A;
  const double gamma = -3 * B4 / (256 * A4) + B2 * C / (16 * A3) - B * D / (4 * A2) + E / A;
  const double alpha2 = alpha * alpha, alpha3 = alpha2 * alpha;
  const pf_cxd P = pf_cx(-alpha2 / 12 - gamma, 0.0);
  const pf_cxd Q = pf_cx(-alpha3 / 108 + alpha * gamma / 3 - beta * beta / 8, 0.0);
  const pf_cxd disc = pf_cx_add(pf_cx_divs(pf_cx_pow_real_int(Q.re, 2), 4.0), pf_cx_divs(pf_cx_pow_real_int(P.re, 3), 27.0));
  const pf_cxd R = pf_cx_add(pf_cx_divs(pf_cx_neg(Q), 2.0), pf_cx_sqrt(disc));
  const pf_cxd U = pf_cx_pow(R, 1.0 / 3.0, true);
  const double a56 = -5.0 * alpha / 6.0;
  pf_cxd y;
  if (U.re == 0.0) {
    const pf_cxd q3 = pf_cx_pow(Q, 1.0 / 3.0, true);
    y = pf_cx(a56 - q3.re, -q3.im);
  } else {
    const pf_cxd t = pf_cx_div(P, pf_cx_scale(U, 3.0));
    y = pf_cx_add(pf_cx(a56 - t.re, -t.im), U);
  }
  const pf_cxd y2 = pf_cx_scale(y, 2.0);
  const pf_cxd w = pf_cx_sqrt(pf_cx(alpha + y2.re, y2.im));
  const pf_cxd bw = pf_cx_div(pf_cx(2.0 * beta, 0.0), w);
  const pf_cxd base = pf_cx(3.0 * alpha + y2.re, y2.im);
  const pf_cxd sp = pf_cx_sqrt(pf_cx_neg(pf_cx_add(base, bw)));
  const pf_cxd sm = pf_cx_sqrt(pf_cx_neg(pf_cx_sub(base, bw)));
  const double sh = -B / (4.0 * A);
  roots[0] = sh + 0.5 * (w.re + sp.re);
  roots[1] = sh + 0.5 * (w.re - sp.re);
  roots[2] = sh + 0.5 * (-w.re + sm.re);
  roots[3] = sh + 0.5 * (-w.re - sm.re);
}

typedef pf_p3p_d3 pf_d3;
__device__ __forceinline__ pf_d3 pf_d3m(double x, double y, double z) { pf_d3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ pf_d3 pf_d3_sub(pf_d3 a, pf_d3 b) { return pf_d3m(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ pf_d3 pf_d3_cross(pf_d3 a, pf_d3 b)
{
  return pf_d3m(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ double pf_d3_dot(pf_d3 a, pf_d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ double pf_d3_norm(pf_d3 a) { return sqrt(pf_d3_dot(a, a)); }
__device__ __forceinline__ pf_d3 pf_d3_divs(pf_d3 a, double s) { return pf_d3m(a.x / s, a.y / s, a.z / s); }

// One P3P problem = a camera part (depends on the three bearing vectors only), a world part (depends on the
// three LEDs and on whether the camera part swapped the first two correspondences) and the quartic that joins
// them.  The vote shares the camera part between all LED permutations of a detection triple and reads the world
// parts from a table the prologue fills; the operations and their order are those of P3P::computePoses either way.
struct pf_p3p_cam {
  pf_d3 e1, e2, e3;     // rows of T (intermediate camera frame)
  double phi1, phi2, b;
  int swapped;          // :104-125 took the branch f3(2) > 0
};

__device__ __forceinline__ void pf_p3p_cam_frame(pf_d3 f1, pf_d3 f2, pf_d3 f3in, pf_p3p_cam& Cc)
{
  pf_d3 e1 = f1, e3 = pf_d3_cross(f1, f2);
  e3 = pf_d3_divs(e3, pf_d3_norm(e3));
  pf_d3 e2 = pf_d3_cross(e3, e1);
  pf_d3 f3 = pf_d3m(pf_d3_dot(e1, f3in), pf_d3_dot(e2, f3in), pf_d3_dot(e3, f3in));
  Cc.swapped = 0;
  if (f3.z > 0.0) {   // :104-125: swap the first two correspondences so that theta lies in [0, pi]
    const pf_d3 t = f1; f1 = f2; f2 = t;
    e1 = f1; e3 = pf_d3_cross(f1, f2);
    e3 = pf_d3_divs(e3, pf_d3_norm(e3));
    e2 = pf_d3_cross(e3, e1);
    f3 = pf_d3m(pf_d3_dot(e1, f3in), pf_d3_dot(e2, f3in), pf_d3_dot(e3, f3in));
    Cc.swapped = 1;
  }
  Cc.e1 = e1; Cc.e2 = e2; Cc.e3 = e3;
  Cc.phi1 = f3.x / f3.z; Cc.phi2 = f3.y / f3.z;
  const double cos_beta = pf_d3_dot(f1, f2);
  const double b = 1 / (1 - cos_beta * cos_beta) - 1;
  Cc.b = cos_beta < 0 ? -sqrt(b) : sqrt(b);
}

// :78-84, on the world points in the caller's order
__device__ __forceinline__ bool pf_p3p_colinear(pf_d3 P1, pf_d3 P2, pf_d3 P3)
{
  return pf_d3_norm(pf_d3_cross(pf_d3_sub(P2, P1), pf_d3_sub(P3, P1))) == 0.0;
}

// world part for the LEDs (P1, P2, P3) in the order the solver uses them (already swapped if the camera part says so)
__device__ __forceinline__ void pf_p3p_world_frame(pf_d3 P1, pf_d3 P2, pf_d3 P3, pf_p3p_world& W)
{
  pf_d3 n1 = pf_d3_sub(P2, P1);
  n1 = pf_d3_divs(n1, pf_d3_norm(n1));
  pf_d3 n3 = pf_d3_cross(n1, pf_d3_sub(P3, P1));
  n3 = pf_d3_divs(n3, pf_d3_norm(n3));
  const pf_d3 n2 = pf_d3_cross(n3, n1);
  const pf_d3 d3 = pf_d3_sub(P3, P1);
  W.n1 = n1; W.n2 = n2; W.n3 = n3; W.P1 = P1;
  W.p1 = pf_d3_dot(n1, d3); W.p2 = pf_d3_dot(n2, d3);
  W.d12 = pf_d3_norm(pf_d3_sub(P2, P1));
}

// Coefficients a[0..4] of Kneip's quartic in cos(theta) (Kneip, Scaramuzza, Siegwart, CVPR 2011, eq. 11: phi1, phi2 = the
// third bearing vector's ratios, p1, p2 = the third world point in the intermediate world frame, d12 = |P2 - P1|,
// b = cot(beta)) and its roots.  The terms are summed in the order of the published polynomial as the reference
// evaluates it (p3p.cpp:150-173): FP64 addition is not associative and the duplicate test downstream compares doubles
// for exact equality.
__device__ __forceinline__ void pf_p3p_roots(const pf_p3p_cam& Cc, const pf_p3p_world& W, double roots[4])
{
  const double phi1 = Cc.phi1, phi2 = Cc.phi2, b = Cc.b, p1 = W.p1, p2 = W.p2, d12 = W.d12;
  const double phi1_sq = phi1 * phi1, phi2_sq = phi2 * phi2, p1_sq = p1 * p1, p1_cu = p1_sq * p1,
               p1_qu = p1_cu * p1, p2_sq = p2 * p2, p2_cu = p2_sq * p2, p2_qu = p2_cu * p2,
               d12_sq = d12 * d12, b_sq = b * b;
  double a[5];
  a[0] = -phi2_sq * p2_qu - p2_qu * phi1_sq - p2_qu;
  a[1] = 2 * p2_cu * d12 * b + 2 * phi2_sq * p2_cu * d12 * b - 2 * phi2 * p2_cu * phi1 * d12;
  a[2] = -phi2_sq * p2_sq * p1_sq - phi2_sq * p2_sq * d12_sq * b_sq - phi2_sq * p2_sq * d12_sq +
           phi2_sq * p2_qu + p2_qu * phi1_sq + 2 * p1 * p2_sq * d12 +
           2 * phi1 * phi2 * p1 * p2_sq * d12 * b - p2_sq * p1_sq * phi1_sq +
           2 * p1 * p2_sq * phi2_sq * d12 - p2_sq * d12_sq * b_sq - 2 * p1_sq * p2_sq;
  a[3] = 2 * p1_sq * p2 * d12 * b + 2 * phi2 * p2_cu * phi1 * d12 - 2 * phi2_sq * p2_cu * d12 * b -
           2 * p1 * p2 * d12_sq * b;
  a[4] = -2 * phi2 * p2_sq * phi1 * p1 * d12 * b + phi2_sq * p2_sq * d12_sq + 2 * p1_cu * d12 -
           p1_sq * d12_sq + phi2_sq * p2_sq * p1_sq - p1_qu - 2 * phi2_sq * p2_sq * p1 * d12 +
           p2_sq * phi1_sq * p1_sq + phi2_sq * p2_sq * d12_sq * b_sq;
  pf_p3p_quartic(a, roots);
}

// Back-substitution of one root (p3p.cpp:181-233): S = [R | C] column-major 3x4, R = N^T R'^T T, C = P1 + N^T C'
__device__ __forceinline__ void pf_p3p_backsub(const pf_p3p_cam& Cc, const pf_p3p_world& W, double r, double S[12])
{
  // evaluated as the reference writes it: the solution entries feed the exact-equality duplicate rule (:1659-1663)
  const double cot_alpha = (-Cc.phi1 * W.p1 / Cc.phi2 - r * W.p2 + W.d12 * Cc.b) / (-Cc.phi1 * r * W.p2 / Cc.phi2 + W.p1 - W.d12);
  const double cos_theta = r, sin_theta = sqrt(1 - r * r);
  const double sin_alpha = sqrt(1 / (cot_alpha * cot_alpha + 1));
  double cos_alpha = sqrt(1 - sin_alpha * sin_alpha);
  if (cot_alpha < 0) cos_alpha = -cos_alpha;
  const double k = sin_alpha * Cc.b + cos_alpha;
  const double Cx = W.d12 * cos_alpha * k, Cy = cos_theta * W.d12 * sin_alpha * k, Cz = sin_theta * W.d12 * sin_alpha * k;
  S[9] = W.P1.x + (W.n1.x * Cx + W.n2.x * Cy + W.n3.x * Cz);
  S[10] = W.P1.y + (W.n1.y * Cx + W.n2.y * Cy + W.n3.y * Cz);
  S[11] = W.P1.z + (W.n1.z * Cx + W.n2.z * Cy + W.n3.z * Cz);
  // R' as written at :205-214; M1 = N^T R'^T : M1(r,c) = sum_k N(k,r) R'(c,k)
  const double Rp[3][3] = {{-cos_alpha, -sin_alpha * cos_theta, -sin_alpha * sin_theta},
                           {sin_alpha, -cos_alpha * cos_theta, -cos_alpha * sin_theta},
                           {0.0, -sin_theta, cos_theta}};
  const double Nm[3][3] = {{W.n1.x, W.n1.y, W.n1.z}, {W.n2.x, W.n2.y, W.n2.z}, {W.n3.x, W.n3.y, W.n3.z}};
  const double Tm[3][3] = {{Cc.e1.x, Cc.e1.y, Cc.e1.z}, {Cc.e2.x, Cc.e2.y, Cc.e2.z}, {Cc.e3.x, Cc.e3.y, Cc.e3.z}};
#pragma unroll
  for (int rr = 0; rr < 3; ++rr) {
    double m[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) m[c] = (Nm[0][rr] * Rp[c][0] + Nm[1][rr] * Rp[c][1]) + Nm[2][rr] * Rp[c][2];
#pragma unroll
    for (int c = 0; c < 3; ++c) S[3 * c + rr] = (m[0] * Tm[0][c] + m[1] * Tm[1][c]) + m[2] * Tm[2][c];
  }
}

// ---- batched solver: one problem per thread (checkCorrespondences' use, pose_estimator.cpp:1312-1501, and tests) ----
// fv, wp: [n][9] column-major 3x3 (columns = bearing vectors / world points); sol: [n][4][12]; rc: [n] (0 or -1)
__global__ void __launch_bounds__(PF_P3P_BLOCK) pf_p3p_solve_kernel(int n, const double* fv, const double* wp, double* sol, int* rc)
{
  const int i = blockIdx.x * PF_P3P_BLOCK + threadIdx.x;
  if (i >= n) return;
  const double* f = fv + 9 * (size_t)i;
  const double* w = wp + 9 * (size_t)i;
  const pf_d3 P1 = pf_d3m(w[0], w[1], w[2]), P2 = pf_d3m(w[3], w[4], w[5]), P3 = pf_d3m(w[6], w[7], w[8]);
  pf_p3p_cam Cc;
  pf_p3p_world W;
  pf_p3p_cam_frame(pf_d3m(f[0], f[1], f[2]), pf_d3m(f[3], f[4], f[5]), pf_d3m(f[6], f[7], f[8]), Cc);
  if (Cc.swapped) pf_p3p_world_frame(P2, P1, P3, W); else pf_p3p_world_frame(P1, P2, P3, W);
  const bool ok = !pf_p3p_colinear(P1, P2, P3);
  rc[i] = ok ? 0 : -1;
  if (!ok) return;
  double roots[4];
  pf_p3p_roots(Cc, W, roots);
#pragma unroll 1
  for (int k = 0; k < 4; ++k) {
    double S[12];
    pf_p3p_backsub(Cc, W, k == 0 ? roots[0] : k == 1 ? roots[1] : k == 2 ? roots[2] : roots[3], S);
    for (int e = 0; e < 12; ++e) sol[48 * (size_t)i + 12 * k + e] = S[e];
  }
}

__device__ __forceinline__ double pf_sqd2(double ax, double ay, double bx, double by)
{
  const double dx = ax - bx, dy = ay - by;
  return dx * dx + dy * dy;
}

// ---- checkCorrespondences (pose_estimator.cpp:1312-1501): the per-combination body, batched ----------------------
// One thread per (correspondence set, combination of three of its pairs): P3P on the three, back-projection of the
// other pairs' LEDs, calculateSquaredReprojectionErrorAndCertainty (:1087-1132) against their detections, selection
// of the best solution that reaches the certainty threshold (:1412-1420) and its inverse -- the pose the reference
// seeds a particle with (:1437) and averages the marker positions through (:1449).
__global__ void __launch_bounds__(PF_P3P_BLOCK) pf_p3p_check_kernel(pf_p3p_params p, int M, int stride, const unsigned* corr,
                                                                    const int* counts, double certainty_threshold, pf_p3p_check_rec* out)
{
  const long long gid = (long long)blockIdx.x * PF_P3P_BLOCK + threadIdx.x;
  const int m = (int)(gid / stride), c = (int)(gid % stride);
  if (m >= M) return;
  const int n = counts[m];
  if (n < 3 || n > PF_P3P_MAX_L || c >= n * (n - 1) * (n - 2) / 6) return;   // stride >= C(n, 3) is checked on the host
  // lexicographic unranking of (a < b < g) out of n (Combinations::combinationsNoReplacement's order)
  int a = 0, rem = c;
  while (rem >= (n - 1 - a) * (n - 2 - a) / 2) { rem -= (n - 1 - a) * (n - 2 - a) / 2; ++a; }
  int b = a + 1;
  while (rem >= n - 1 - b) { rem -= n - 1 - b; ++b; }
  const int g = b + 1 + rem;
  const unsigned* row = corr + (size_t)m * 2 * PF_P3P_MAX_L;
  pf_p3p_check_rec R;
  R.rc = 0; R.valid = 0; R.best = -1; R.reserved = 0; R.squared_error = INFINITY;
#pragma unroll
  for (int e = 0; e < 16; ++e) R.pose[e] = 0.0;
  bool bad = false;
  for (int l = 0; l < n; ++l) bad |= row[2 * l] < 1u || row[2 * l] > (unsigned)p.L || row[2 * l + 1] < 1u || row[2 * l + 1] > (unsigned)p.D;
  if (bad) { R.rc = -2; out[(size_t)m * stride + c] = R; return; }
  const int tri[3] = {a, b, g};
  pf_d3 f[3], P[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int led = (int)row[2 * tri[k]] - 1, det = (int)row[2 * tri[k] + 1] - 1;
    P[k] = pf_d3m(p.obj[3 * led], p.obj[3 * led + 1], p.obj[3 * led + 2]);
    pf_d3 v = pf_d3m((p.img[2 * det] - p.K[2]) / p.K[0], (p.img[2 * det + 1] - p.K[5]) / p.K[4], 1.0);   // :1072-1085
    f[k] = pf_d3_divs(v, pf_d3_norm(v));
  }
  if (pf_p3p_colinear(P[0], P[1], P[2])) { R.rc = -1; out[(size_t)m * stride + c] = R; return; }
  pf_p3p_cam Cc;
  pf_p3p_world W;
  pf_p3p_cam_frame(f[0], f[1], f[2], Cc);
  if (Cc.swapped) pf_p3p_world_frame(P[1], P[0], P[2], W); else pf_p3p_world_frame(P[0], P[1], P[2], W);
  double roots[4];
  pf_p3p_roots(Cc, W, roots);
  const double tol2 = p.tol * p.tol;
#pragma unroll 1
  for (int j = 0; j < 4; ++j) {
    double S[12];
    pf_p3p_backsub(Cc, W, j == 0 ? roots[0] : j == 1 ? roots[1] : j == 2 ? roots[2] : roots[3], S);
    bool finite = true;
#pragma unroll
    for (int e = 0; e < 12; ++e) finite &= ((S[e] - S[e]) == (S[e] - S[e]));
    if (!finite) continue;
    const double a00 = S[0], a10 = S[1], a20 = S[2], a01 = S[3], a11 = S[4], a21 = S[5], a02 = S[6], a12 = S[7], a22 = S[8];
    const double c00 = a11 * a22 - a12 * a21, c01 = a12 * a20 - a10 * a22, c02 = a10 * a21 - a11 * a20;
    const double det = a00 * c00 + a01 * c01 + a02 * c02;
    const double i00 = c00 / det, i01 = (a02 * a21 - a01 * a22) / det, i02 = (a01 * a12 - a02 * a11) / det;
    const double i10 = c01 / det, i11 = (a00 * a22 - a02 * a20) / det, i12 = (a02 * a10 - a00 * a12) / det;
    const double i20 = c02 / det, i21 = (a01 * a20 - a00 * a21) / det, i22 = (a00 * a11 - a01 * a10) / det;
    const double t0 = -(i00 * S[9] + i01 * S[10] + i02 * S[11]), t1 = -(i10 * S[9] + i11 * S[10] + i12 * S[11]),
                 t2 = -(i20 * S[9] + i21 * S[10] + i22 * S[11]);
    double d[PF_P3P_MAX_L];
    int n_un = 0;
    for (int l = 0; l < n; ++l) {
      if (l == a || l == b || l == g) continue;
      const int led = (int)row[2 * l] - 1, dt = (int)row[2 * l + 1] - 1;
      const double X = p.obj[3 * led], Y = p.obj[3 * led + 1], Z = p.obj[3 * led + 2];
      const double xc = i00 * X + i01 * Y + i02 * Z + t0, yc = i10 * X + i11 * Y + i12 * Z + t1, zc = i20 * X + i21 * Y + i22 * Z + t2;
      const double q0 = p.K[0] * xc + p.K[1] * yc + p.K[2] * zc, q1 = p.K[3] * xc + p.K[4] * yc + p.K[5] * zc,
                   q2 = p.K[6] * xc + p.K[7] * yc + p.K[8] * zc;
      d[n_un++] = pf_sqd2(p.img[2 * dt], p.img[2 * dt + 1], q0 / q2, q1 / q2);
    }
    double se = 0.0;
    int num = 0;
    for (int r = 0; r < n_un; ++r) {   // minCoeff / set to INF until the minimum exceeds tol^2 (:1112-1126)
      int best = 0;
      for (int i = 1; i < n_un; ++i) if (d[i] < d[best]) best = i;
      if (!(d[best] <= tol2)) break;
      se += d[best]; ++num; d[best] = INFINITY;
    }
    const double certainty = (double)num / (double)n_un;   // 0/0 = NaN without unused pairs: never reaches a threshold
    if (certainty >= certainty_threshold) {
      R.valid = 1;
      if (se < R.squared_error) {
        R.squared_error = se; R.best = j;
        R.pose[0] = i00; R.pose[1] = i10; R.pose[2] = i20; R.pose[3] = 0.0;
        R.pose[4] = i01; R.pose[5] = i11; R.pose[6] = i21; R.pose[7] = 0.0;
        R.pose[8] = i02; R.pose[9] = i12; R.pose[10] = i22; R.pose[11] = 0.0;
        R.pose[12] = t0; R.pose[13] = t1; R.pose[14] = t2; R.pose[15] = 1.0;
      }
    }
  }
  out[(size_t)m * stride + c] = R;
}

// LED permutation number -> (j0, j1, j2), all different, lexicographic in (j0, j1, j2)
__device__ __forceinline__ void pf_p3p_perm(int pr, int L, int& j0, int& j1, int& j2)
{
  j0 = pr / ((L - 1) * (L - 2));
  pr -= j0 * (L - 1) * (L - 2);
  j1 = pr / (L - 2); j2 = pr % (L - 2);
  j1 += j1 >= j0 ? 1 : 0;
  const int lo = j0 < j1 ? j0 : j1, hi = j0 < j1 ? j1 : j0;
  j2 += j2 >= lo ? 1 : 0;
  j2 += j2 >= hi ? 1 : 0;
}

// ---- prologue: bearing vectors, sorted detections, world-frame table, cleared outputs -----------------------------
__global__ void pf_p3p_prep_kernel(pf_p3p_params p)
{
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsz = gridDim.x * blockDim.x;
  for (int i = gtid; i < p.D; i += gsz) {
    pf_d3 v = pf_d3m((p.img[2 * i] - p.K[2]) / p.K[0], (p.img[2 * i + 1] - p.K[5]) / p.K[4], 1.0);
    v = pf_d3_divs(v, pf_d3_norm(v));
    p.vec[3 * i] = v.x; p.vec[3 * i + 1] = v.y; p.vec[3 * i + 2] = v.z;
  }
  // detections sorted by x (rank sort; a NaN column sorts last and is never inside a window)
  for (int i = gtid; i < p.D; i += gsz) {
    const double xi = p.img[2 * i], ki = xi != xi ? INFINITY : xi;
    int rank = 0;
    for (int j = 0; j < p.D; ++j) {
      const double xj = p.img[2 * j], kj = xj != xj ? INFINITY : xj;
      rank += (kj < ki || (kj == ki && j < i)) ? 1 : 0;
    }
    p.simg[2 * rank] = xi; p.simg[2 * rank + 1] = p.img[2 * i + 1];
    p.sidx[rank] = i;
  }
  // world part of every LED permutation, as given and with the first two LEDs swapped (:122-124)
  const int nperm = p.L * (p.L - 1) * (p.L - 2);
  for (int e = gtid; e < 2 * nperm; e += gsz) {
    int j0, j1, j2;
    pf_p3p_perm(e >> 1, p.L, j0, j1, j2);
    const pf_d3 P1 = pf_d3m(p.obj[3 * j0], p.obj[3 * j0 + 1], p.obj[3 * j0 + 2]),
                P2 = pf_d3m(p.obj[3 * j1], p.obj[3 * j1 + 1], p.obj[3 * j1 + 2]),
                P3 = pf_d3m(p.obj[3 * j2], p.obj[3 * j2 + 1], p.obj[3 * j2 + 2]);
    pf_p3p_world W;
    if (e & 1) pf_p3p_world_frame(P2, P1, P3, W); else pf_p3p_world_frame(P1, P2, P3, W);
    W.colinear = pf_p3p_colinear(P1, P2, P3) ? 1.0 : 0.0;
    p.wtab[e] = W;
  }
  for (int i = gtid; i < p.D * p.L; i += gsz) p.hist[i] = 0u;
  if (gtid < 4) p.stats[gtid] = 0ull;
}

// number of triples whose first element is < a, out of D elements: C(D,3) - C(D-a,3)
__device__ __forceinline__ long long pf_c3(long long m) { return m * (m - 1) * (m - 2) / 6; }
__device__ __forceinline__ long long pf_c2(long long m) { return m * (m - 1) / 2; }

// ---- the vote (pose_estimator.cpp:1540-1717) ---------------------------------------------------------------------
// One thread owns one detection triple (i0 < i1 < i2, lexicographic rank = its number) and a slice of the LED
// permutations (p.nchunk slices per triple, so that small problems still fill the GPU): pre-filters and camera part
// once, then per permutation the quartic, four back-substitutions, back-projection and votes.
// dynamic shared memory: sxy [D] double2 | xb [2] double | hist [D*L] u32 | sid [D] int | first [PF_P3P_NB + 1] int | sprev [12][threads] double
// The detections near a projected LED are found through a bucket table over x (first[b] = the first sorted detection
// whose bucket is >= b; bucket = a monotone function of x, so nothing left of first[bucket(x_lo)] can be >= x_lo)
// instead of a binary search per projected LED: one shared-memory load in place of eight dependent ones.
#define PF_P3P_NB 256
__device__ __forceinline__ int pf_p3p_bucket(double x, double x0, double inv_w)
{
  const double t = (x - x0) * inv_w;
  if (x != x) return PF_P3P_NB;                               // NaN sorts last
  return t >= (double)PF_P3P_NB ? PF_P3P_NB : (t > 0.0 ? (int)t : 0);
}
// LT: the LED count as a compile-time constant (5: the unused LEDs of a permutation and their projections stay in
// registers), or 0 for any count up to PF_P3P_MAX_L (local arrays)
// MODE 0: everything in one kernel.  MODE 1 / 2: the same work as two launches -- 1 = up to the four roots of every
// (triple, permutation), written to rbuf [step][thread] (32 B each); 2 = from those roots on (back-substitution,
// back-projection, votes).  The single kernel is 80 KB of code and waits for instruction fetch more than for anything
// else (`no_instruction` 5.7 stall cycles per issued instruction); the second half alone fits the 32 KB instruction
// cache.  gid0 / nthr: the slab of threads this launch covers (rbuf holds one slab).
template <int MINB, int LT, int MODE>
__global__ void __launch_bounds__(PF_P3P_BLOCK, MINB) pf_p3p_hist_kernel(pf_p3p_params p, double2* rbuf, long long gid0, long long nthr)
{
  constexpr int NU = LT ? LT - 3 : PF_P3P_MAX_L - 3;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int D = p.D, L = LT ? LT : p.L;
  double2* sxy = reinterpret_cast<double2*>(smem_raw);
  double* xb = reinterpret_cast<double*>(sxy + D);
  unsigned* sh = reinterpret_cast<unsigned*>(xb + 2);
  int* sid = reinterpret_cast<int*>(sh + D * L);
  int* first = sid + D;
  const int tid = threadIdx.x;
  // the previous solution of this thread's problem (duplicate rule, :1659-1663): [12][threads], behind the tables
  double* sprev = reinterpret_cast<double*>(smem_raw + ((size_t)D * 16 + 16 + (size_t)D * L * 4 + (size_t)D * 4 + (size_t)(PF_P3P_NB + 1) * 4 + 15) / 16 * 16) + tid;
  for (int i = tid; i < D; i += PF_P3P_BLOCK) {
    sxy[i] = make_double2(p.simg[2 * i], p.simg[2 * i + 1]);
    sid[i] = p.sidx[i];
  }
  for (int i = tid; i < D * L; i += PF_P3P_BLOCK) sh[i] = 0u;
  __syncthreads();
  if (MODE != 1 && tid == 0) {         // bucket grid over [smallest x, largest finite x]
    int e = D - 1;
    while (e > 0 && !(fabs(sxy[e].x) < INFINITY)) --e;
    const double span = sxy[e].x - sxy[0].x;
    xb[0] = sxy[0].x;
    xb[1] = (span > 0.0 && span < INFINITY) ? (double)PF_P3P_NB / span : 0.0;
  }
  __syncthreads();
  const double bx0 = MODE != 1 ? xb[0] : 0.0, binv = MODE != 1 ? xb[1] : 0.0;
  for (int b = tid; MODE != 1 && b <= PF_P3P_NB; b += PF_P3P_BLOCK) {
    int lo = 0, hi = D;
    while (lo < hi) {                  // first sorted detection whose bucket is >= b
      const int mid = (lo + hi) >> 1;
      if (pf_p3p_bucket(sxy[mid].x, bx0, binv) < b) lo = mid + 1; else hi = mid;
    }
    first[b] = lo;
  }
  __syncthreads();

  const long long n_triples_all = pf_c3(D);
  const long long lgid = (long long)blockIdx.x * PF_P3P_BLOCK + tid;     // thread of this launch (slab)
  const long long gid = gid0 + lgid;
  const long long triple = gid / p.nchunk;
  const int chunk = (int)(gid % p.nchunk);
  const bool in_range = triple < n_triples_all && lgid < nthr;
  // unrank: i0 = largest a with C(D,3) - C(D-a,3) <= triple; then the pair (i1, i2) inside the D-1-i0 that follow
  int i0 = 0, i1 = 1, i2 = 2;
  if (in_range) {
    const long long rest = n_triples_all - triple;            // triples from this one to the end, >= 1
    int m = (int)cbrt(6.0 * (double)rest) + 2;                // smallest m with C(m,3) >= rest, then fix up
    if (m > D) m = D;
    while (m > 3 && pf_c3(m - 1) >= rest) --m;
    while (pf_c3(m) < rest) ++m;
    i0 = D - m;
    const long long r2 = triple - (n_triples_all - pf_c3(m));  // rank of (i1, i2) among pairs of the m-1 followers
    const long long rest2 = pf_c2(m - 1) - r2;                 // >= 1
    int q = (int)sqrt(2.0 * (double)rest2) + 2;                // smallest q with C(q,2) >= rest2
    if (q > m - 1) q = m - 1;
    while (q > 2 && pf_c2(q - 1) >= rest2) --q;
    while (pf_c2(q) < rest2) ++q;
    i1 = D - q;
    i2 = i1 + 1 + (int)(r2 - (pf_c2(m - 1) - pf_c2(q)));
  }

  const double thresh = 10000.0 * 100.0;   // threshDist, threshDist2 (:1558-1559)
  const double d1x = p.img[2 * i0], d1y = p.img[2 * i0 + 1], d2x = p.img[2 * i1], d2y = p.img[2 * i1 + 1],
               d3x = p.img[2 * i2], d3y = p.img[2 * i2 + 1];
  const double mx = (d1x + d2x + d3x) / 3, my = (d1y + d2y + d3y) / 3;
  // pre-filters of the triple (:1561-1581)
  bool pass = in_range && !(pf_sqd2(d1x, d1y, d2x, d2y) > thresh) && !(pf_sqd2(d1x, d1y, d3x, d3y) > thresh) &&
              !(pf_sqd2(d2x, d2y, d3x, d3y) > thresh);
  if (pass) {
    int cd = 0;
    for (int q = 0; q < D; ++q) cd += pf_sqd2(mx, my, sxy[q].x, sxy[q].y) < thresh ? 1 : 0;
    pass = cd >= 5;
  }
  unsigned n_triples = (pass && chunk == 0 && MODE != 1) ? 1u : 0u, n_solved = 0, n_proj = 0, n_vote = 0;

  pf_p3p_cam Cc;
  pf_p3p_cam_frame(pf_d3m(p.vec[3 * i0], p.vec[3 * i0 + 1], p.vec[3 * i0 + 2]),
                   pf_d3m(p.vec[3 * i1], p.vec[3 * i1 + 1], p.vec[3 * i1 + 2]),
                   pf_d3m(p.vec[3 * i2], p.vec[3 * i2 + 1], p.vec[3 * i2 + 2]), Cc);
  const int nperm = L * (L - 1) * (L - 2);
  const int per = (nperm + p.nchunk - 1) / p.nchunk;
  const int pr_begin = chunk * per, pr_end = min(nperm, pr_begin + per);
  const double tol = p.tol;
  const double reach = tol + 1e-6;   // column window; every candidate is then tested exactly
  const int n_un = LT ? NU : L - 3;

  // The loops are written so that a warp stays converged: every thread runs `per` steps, dead threads skip the
  // bodies, and the warp re-joins after every solution.
#pragma unroll 1
  for (int step = 0; step < per; ++step) {
    const int pr = pr_begin + step;
    bool live = pass && pr < pr_end;
    int j0, j1, j2;
    pf_p3p_perm(live ? pr : 0, L, j0, j1, j2);
    // the world part is read through the table (one 128-byte line, L1-resident) where it is used, not copied: 32 fewer
    // live registers in a kernel that spills
    const pf_p3p_world& W = p.wtab[2 * (live ? pr : 0) + Cc.swapped];
    live = live && W.colinear == 0.0;
    n_solved += (live && MODE != 1) ? 1u : 0u;
    double r0 = 0, r1 = 0, r2 = 0, r3 = 0;
    if (MODE == 2) {
      if (live) {
        const double2 a = rbuf[2 * ((long long)step * nthr + lgid)], b = rbuf[2 * ((long long)step * nthr + lgid) + 1];
        r0 = a.x; r1 = a.y; r2 = b.x; r3 = b.y;
      }
    } else if (live) {
      double roots[4];
      pf_p3p_roots(Cc, W, roots);
      r0 = roots[0]; r1 = roots[1]; r2 = roots[2]; r3 = roots[3];
    }
    if (MODE == 1) {
      if (live) {
        rbuf[2 * ((long long)step * nthr + lgid)] = make_double2(r0, r1);
        rbuf[2 * ((long long)step * nthr + lgid) + 1] = make_double2(r2, r3);
      }
      continue;
    }
    __syncwarp();
#pragma unroll 1
    for (int k = 0; k < 4; ++k) {
      double S[12];
      bool use = false;
      asm volatile("" ::: "memory");     // table entries are re-read per solution rather than kept across the loop
      if (live) {
        pf_p3p_backsub(Cc, W, k == 0 ? r0 : k == 1 ? r1 : k == 2 ? r2 : r3, S);
        bool skip = false;
        if (k > 0) {   // :1659-1663: a solution is dropped when ANY entry equals the previous solution's
#pragma unroll
          for (int e = 0; e < 12; ++e) skip |= (S[e] - sprev[e * PF_P3P_BLOCK] == 0.0);
        }
        bool finite = true;   // isFinite(H) (:1670)
#pragma unroll
        for (int e = 0; e < 12; ++e) { finite &= ((S[e] - S[e]) == (S[e] - S[e])); sprev[e * PF_P3P_BLOCK] = S[e]; }
        use = !skip && finite;
      }
      __syncwarp();
      if (use) {
        ++n_proj;
        // inverse of [R C; 0 1] by cofactors: Ai = adj(R) / det(R) (one reciprocal), ti = -Ai C
        const double a00 = S[0], a10 = S[1], a20 = S[2], a01 = S[3], a11 = S[4], a21 = S[5], a02 = S[6], a12 = S[7], a22 = S[8];
        const double c00 = a11 * a22 - a12 * a21, c01 = a12 * a20 - a10 * a22, c02 = a10 * a21 - a11 * a20;
        const double idet = 1.0 / (a00 * c00 + a01 * c01 + a02 * c02);
        const double i00 = c00 * idet, i01 = (a02 * a21 - a01 * a22) * idet, i02 = (a01 * a12 - a02 * a11) * idet;
        const double i10 = c01 * idet, i11 = (a00 * a22 - a02 * a20) * idet, i12 = (a02 * a10 - a00 * a12) * idet;
        const double i20 = c02 * idet, i21 = (a01 * a20 - a00 * a21) * idet, i22 = (a00 * a11 - a01 * a10) * idet;
        const double t0 = -(i00 * S[9] + i01 * S[10] + i02 * S[11]), t1 = -(i10 * S[9] + i11 * S[10] + i12 * S[11]),
                     t2 = -(i20 * S[9] + i21 * S[10] + i22 * S[11]);
        // project the unused LEDs (ascending LED number, :1637-1656) with the full K (project2d :1017-1034)
        double pu[NU], pv[NU];
        int pl[NU];
        auto project = [&](int ll, double& u, double& v) {
          const double X = p.obj[3 * ll], Y = p.obj[3 * ll + 1], Z = p.obj[3 * ll + 2];
          const double xc = i00 * X + i01 * Y + i02 * Z + t0, yc = i10 * X + i11 * Y + i12 * Z + t1,
                       zc = i20 * X + i21 * Y + i22 * Z + t2;
          const double q0 = p.K[0] * xc + p.K[1] * yc + p.K[2] * zc, q1 = p.K[3] * xc + p.K[4] * yc + p.K[5] * zc,
                       q2 = p.K[6] * xc + p.K[7] * yc + p.K[8] * zc;
          const double iq2 = 1.0 / q2;
          u = q0 * iq2; v = q1 * iq2;
        };
        if (LT) {
          unsigned mk = ((1u << LT) - 1u) ^ ((1u << j0) | (1u << j1) | (1u << j2));
#pragma unroll
          for (int a = 0; a < NU; ++a) {
            pl[a] = __ffs(mk) - 1;
            mk &= mk - 1u;
            project(pl[a], pu[a], pv[a]);
          }
        } else {
          int m = 0;
          for (int ll = 0; ll < L; ++ll) {
            if (ll == j0 || ll == j1 || ll == j2) continue;
            project(ll, pu[m], pv[m]);
            pl[m] = ll;
            ++m;
          }
        }
        // votes (:1680-1712): every unused detection close to the triple's centroid whose nearest projected LED is
        // closer than tol votes for (detection, that LED); if any did, the triple's own three pairs vote too
        unsigned votes = 0;
#pragma unroll
        for (int a = 0; a < NU; ++a) {
          if (!LT && a >= n_un) break;
          const double u = pu[a], xlo = u - reach, xhi = u + reach;
          for (int q = first[pf_p3p_bucket(xlo, bx0, binv)]; q < D; ++q) {   // sorted detections with xlo <= x <= xhi
            const double bx = sxy[q].x, by = sxy[q].y;
            if (bx < xlo) continue;
            if (!(bx <= xhi)) break;
            if (!(fabs(by - pv[a]) <= reach)) continue;          // farther than tol from this LED: cannot vote for it
            const int det_i = sid[q];
            if (det_i == i0 || det_i == i1 || det_i == i2) continue;
            if (!(pf_sqd2(mx, my, bx, by) < thresh)) continue;   // :1598
            double best = INFINITY;
            int bj = -1;
#pragma unroll
            for (int c = 0; c < NU; ++c) {   // calculateMinDistancesAndPairs (:2093-2137): first strict minimum
              if (!LT && c >= n_un) break;
              const double d2 = pf_sqd2(bx, by, pu[c], pv[c]);
              if (d2 < best) { best = d2; bj = c; }
            }
            if (bj == a && best <= reach * reach && sqrt(best) < tol) {
              atomicAdd(&sh[det_i * L + pl[a]], 1u);
              ++votes;
            }
          }
        }
        if (votes) {
          ++n_vote;
          atomicAdd(&sh[i0 * L + j0], 1u);
          atomicAdd(&sh[i1 * L + j1], 1u);
          atomicAdd(&sh[i2 * L + j2], 1u);
        }
      }
      __syncwarp();
    }
  }
  if (MODE == 1) return;
  __syncthreads();
  for (int i = tid; i < D * L; i += PF_P3P_BLOCK) {
    const unsigned v = sh[i];
    if (v) atomicAdd(&p.hist[i], v);
  }
  // counters: warp sums, one atomic per warp and counter
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    n_triples += __shfl_xor_sync(0xffffffffu, n_triples, o);
    n_solved += __shfl_xor_sync(0xffffffffu, n_solved, o);
    n_proj += __shfl_xor_sync(0xffffffffu, n_proj, o);
    n_vote += __shfl_xor_sync(0xffffffffu, n_vote, o);
  }
  if ((tid & 31) == 0) {
    if (n_triples) atomicAdd(&p.stats[0], (unsigned long long)n_triples);
    if (n_solved) atomicAdd(&p.stats[1], (unsigned long long)n_solved);
    if (n_proj) atomicAdd(&p.stats[2], (unsigned long long)n_proj);
    if (n_vote) atomicAdd(&p.stats[3], (unsigned long long)n_vote);
  }
}

size_t pf_p3p_smem_bytes(int D, int L)
{
  return ((size_t)D * 16 + 16 + (size_t)D * L * 4 + (size_t)D * 4 + (size_t)(PF_P3P_NB + 1) * 4 + 15) / 16 * 16 + (size_t)12 * PF_P3P_BLOCK * 8;
}

template <int LT, int MODE>
static void (*pf_p3p_pick(int occ))(pf_p3p_params, double2*, long long, long long)
{
  return occ >= 8 ? pf_p3p_hist_kernel<8, LT, MODE> : occ == 6 ? pf_p3p_hist_kernel<6, LT, MODE> : occ == 5 ? pf_p3p_hist_kernel<5, LT, MODE>
       : occ == 4 ? pf_p3p_hist_kernel<4, LT, MODE> : pf_p3p_hist_kernel<3, LT, MODE>;
}

size_t pf_p3p_roots_bytes_per_thread(int L, int nchunk)
{
  const int nperm = L * (L - 1) * (L - 2);
  return (size_t)((nperm + nchunk - 1) / nchunk) * 32;
}

// rbuf / rbuf_bytes: scratch for the roots (pf_p3p_roots_bytes_per_thread per thread of a slab), or nullptr / 0: one kernel
cudaError_t pf_p3p_launch_hist(const pf_p3p_params& p, cudaStream_t stream, double2* rbuf, size_t rbuf_bytes)
{
  const size_t smem = pf_p3p_smem_bytes(p.D, p.L);
  // resident CTAs per SM = register budget 168 / 128 / 96 / 80 / 64: the kernel waits on FP64 dependency chains and on
  // instruction fetch (80 KB of code), so more warps beat fewer spills (measured on one B200, D = 205: 44.8 / 39.4 / 38.7 /
  // 36.5 / 36.1 ms for 3 / 4 / 5 / 6 / 8; D = 45: 0.51 / 0.45 / 0.48 / 0.52 / 0.41 ms).  PFMPE_P3P_OCC overrides (tuning aid).
  // Tried and dropped: one out-of-line copy of the FP64 division / square root for the ~45 sites of the inner loop (code
  // 5000 -> 4536 instructions, but 36.1 -> 38.2 ms: the calls cost more than the fetches they save).
  static int occ_env = -1, occ_a_env = -1;
  if (occ_env < 0) { const char* e = getenv("PFMPE_P3P_OCC"); occ_env = e ? atoi(e) : 0; }
  if (occ_a_env < 0) { const char* e = getenv("PFMPE_P3P_OCC_ROOTS"); occ_a_env = e ? atoi(e) : 0; }
  const int occ = occ_env ? occ_env : 8;
  const int occ_a = occ_a_env ? occ_a_env : occ;
  const bool l5 = p.L == 5;
  const long long triples = (long long)p.D * (p.D - 1) * (p.D - 2) / 6;
  const long long threads = triples * p.nchunk;
  pf_p3p_prep_kernel<<<(p.D + 127) / 128 + 2, 128, 0, stream>>>(p);
  const size_t per_thread = pf_p3p_roots_bytes_per_thread(p.L, p.nchunk);
  long long slab = rbuf ? (long long)(rbuf_bytes / per_thread) : 0;
  if (slab >= threads) slab = threads;                                 // everything in one slab
  else if (slab >= 8 * 148 * PF_P3P_BLOCK) {                           // equal slabs, whole CTAs
    const long long cap = slab - slab % PF_P3P_BLOCK, n_slabs = (threads + cap - 1) / cap;
    slab = (threads + n_slabs - 1) / n_slabs;
    slab = (slab + PF_P3P_BLOCK - 1) / PF_P3P_BLOCK * PF_P3P_BLOCK;
    if (slab > cap) slab = cap;
  } else slab = 0;                                                     // less than a wave per launch: not worth it
  auto prepare = [&](void (*k)(pf_p3p_params, double2*, long long, long long)) -> cudaError_t {
    return smem > 48 * 1024 ? cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) : cudaSuccess;
  };
  if (slab == 0) {
    auto kern = l5 ? pf_p3p_pick<5, 0>(occ) : pf_p3p_pick<0, 0>(occ);
    cudaError_t e = prepare(kern);
    if (e != cudaSuccess) return e;
    kern<<<(unsigned)((threads + PF_P3P_BLOCK - 1) / PF_P3P_BLOCK), PF_P3P_BLOCK, smem, stream>>>(p, nullptr, 0, threads);
    return cudaGetLastError();
  }
  auto ka = l5 ? pf_p3p_pick<5, 1>(occ_a) : pf_p3p_pick<0, 1>(occ_a);
  auto kb = l5 ? pf_p3p_pick<5, 2>(occ) : pf_p3p_pick<0, 2>(occ);
  cudaError_t e = prepare(ka);
  if (e == cudaSuccess) e = prepare(kb);
  if (e != cudaSuccess) return e;
  for (long long g0 = 0; g0 < threads; g0 += slab) {
    const long long n = threads - g0 < slab ? threads - g0 : slab;
    const unsigned grid = (unsigned)((n + PF_P3P_BLOCK - 1) / PF_P3P_BLOCK);
    ka<<<grid, PF_P3P_BLOCK, smem, stream>>>(p, rbuf, g0, n);
    kb<<<grid, PF_P3P_BLOCK, smem, stream>>>(p, rbuf, g0, n);
  }
  return cudaGetLastError();
}

cudaError_t pf_p3p_launch_check(const pf_p3p_params& p, int M, int stride, const unsigned* corr, const int* counts,
                                double certainty_threshold, pf_p3p_check_rec* out, cudaStream_t stream)
{
  const long long threads = (long long)M * stride;
  pf_p3p_check_kernel<<<(unsigned)((threads + PF_P3P_BLOCK - 1) / PF_P3P_BLOCK), PF_P3P_BLOCK, 0, stream>>>(p, M, stride, corr, counts,
                                                                                                        certainty_threshold, out);
  return cudaGetLastError();
}

cudaError_t pf_p3p_launch_solve(int n, const double* fv, const double* wp, double* sol, int* rc, cudaStream_t stream)
{
  pf_p3p_solve_kernel<<<(n + PF_P3P_BLOCK - 1) / PF_P3P_BLOCK, PF_P3P_BLOCK, 0, stream>>>(n, fv, wp, sol, rc);
  return cudaGetLastError();
}

