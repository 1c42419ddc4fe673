// pf_p3p.cu -- kernels of K9; see pf_p3p.cuh.  Built with -fmad=false.
#include "pf_p3p.cuh"

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
__device__ __noinline__ pf_cxd pf_cx_pow(pf_cxd z, double y)
{
  if (z.im == 0.0 && z.re > 0.0) return pf_cx(pow(z.re, y), 0.0);
  const double lr = log(hypot(z.re, z.im)), th = atan2(z.im, z.re);
  const double rho = exp(y * lr);
  double sn, cs;
  sincos(y * th, &sn, &cs);
  return pf_cx(rho * cs, rho * sn);
}

// P3P::solveQuartic (p3p.cpp:238-292): Ferrari's closed form, real parts of the four roots
__device__ __noinline__ void pf_p3p_quartic(const double f[5], double roots[4])
{
  const double A = f[0], B = f[1], C = f[2], D = f[3], E = f[4];
  const double A2 = A * A, B2 = B * B, A3 = A2 * A, B3 = B2 * B, A4 = A3 * A, B4 = B3 * B;
  const double alpha = -3 * B2 / (8 * A2) + C / A;
  const double beta = B3 / (8 * A3) - B * C / (2 * A2) + D / A;
  const double gamma = -3 * B4 / (256 * A4) + B2 * C / (16 * A3) - B * D / (4 * A2) + E / A;
  const double alpha2 = alpha * alpha, alpha3 = alpha2 * alpha;
  const pf_cxd P = pf_cx(-alpha2 / 12 - gamma, 0.0);
  const pf_cxd Q = pf_cx(-alpha3 / 108 + alpha * gamma / 3 - beta * beta / 8, 0.0);
  const pf_cxd disc = pf_cx_add(pf_cx_divs(pf_cx_pow(Q, 2.0), 4.0), pf_cx_divs(pf_cx_pow(P, 3.0), 27.0));
  const pf_cxd R = pf_cx_add(pf_cx_divs(pf_cx_neg(Q), 2.0), pf_cx_sqrt(disc));
  const pf_cxd U = pf_cx_pow(R, 1.0 / 3.0);
  const double a56 = -5.0 * alpha / 6.0;
  pf_cxd y;
  if (U.re == 0.0) {
    const pf_cxd q3 = pf_cx_pow(Q, 1.0 / 3.0);
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

struct pf_d3 { double x, y, z; };
__device__ __forceinline__ pf_d3 pf_d3m(double x, double y, double z) { pf_d3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ pf_d3 pf_d3_sub(pf_d3 a, pf_d3 b) { return pf_d3m(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ pf_d3 pf_d3_cross(pf_d3 a, pf_d3 b)
{
  return pf_d3m(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ double pf_d3_dot(pf_d3 a, pf_d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ double pf_d3_norm(pf_d3 a) { return sqrt(pf_d3_dot(a, a)); }
__device__ __forceinline__ pf_d3 pf_d3_divs(pf_d3 a, double s) { return pf_d3m(a.x / s, a.y / s, a.z / s); }

// Frame data of one P3P problem that the four back-substitutions share.
struct pf_p3p_frame {
  pf_d3 e1, e2, e3;     // rows of T (intermediate camera frame)
  pf_d3 n1, n2, n3;     // rows of N (intermediate world frame)
  pf_d3 P1;
  double f_1, f_2, p_1, p_2, d_12, b;
  double roots[4];
};

// P3P::computePoses up to the roots (p3p.cpp:65-179).  Returns false for colinear world points (:78-84).
__device__ __noinline__ bool pf_p3p_setup(pf_d3 f1, pf_d3 f2, pf_d3 f3in, pf_d3 P1, pf_d3 P2, pf_d3 P3, pf_p3p_frame& F)
{
  if (pf_d3_norm(pf_d3_cross(pf_d3_sub(P2, P1), pf_d3_sub(P3, P1))) == 0.0) return false;
  pf_d3 e1 = f1, e3 = pf_d3_cross(f1, f2);
  e3 = pf_d3_divs(e3, pf_d3_norm(e3));
  pf_d3 e2 = pf_d3_cross(e3, e1);
  pf_d3 f3 = pf_d3m(pf_d3_dot(e1, f3in), pf_d3_dot(e2, f3in), pf_d3_dot(e3, f3in));
  if (f3.z > 0.0) {   // :104-125: swap the first two correspondences so that theta lies in [0, pi]
    pf_d3 t = f1; f1 = f2; f2 = t;
    e1 = f1; e3 = pf_d3_cross(f1, f2);
    e3 = pf_d3_divs(e3, pf_d3_norm(e3));
    e2 = pf_d3_cross(e3, e1);
    f3 = pf_d3m(pf_d3_dot(e1, f3in), pf_d3_dot(e2, f3in), pf_d3_dot(e3, f3in));
    t = P1; P1 = P2; P2 = t;
  }
  pf_d3 n1 = pf_d3_sub(P2, P1);
  n1 = pf_d3_divs(n1, pf_d3_norm(n1));
  pf_d3 n3 = pf_d3_cross(n1, pf_d3_sub(P3, P1));
  n3 = pf_d3_divs(n3, pf_d3_norm(n3));
  const pf_d3 n2 = pf_d3_cross(n3, n1);
  const pf_d3 d3 = pf_d3_sub(P3, P1);
  const double p_1 = pf_d3_dot(n1, d3), p_2 = pf_d3_dot(n2, d3);
  const double d_12 = pf_d3_norm(pf_d3_sub(P2, P1));
  const double f_1 = f3.x / f3.z, f_2 = f3.y / f3.z;
  const double cos_beta = pf_d3_dot(f1, f2);
  double b = 1 / (1 - cos_beta * cos_beta) - 1;
  b = cos_beta < 0 ? -sqrt(b) : sqrt(b);

  const double f_1_pw2 = f_1 * f_1, f_2_pw2 = f_2 * f_2, p_1_pw2 = p_1 * p_1, p_1_pw3 = p_1_pw2 * p_1,
               p_1_pw4 = p_1_pw3 * p_1, p_2_pw2 = p_2 * p_2, p_2_pw3 = p_2_pw2 * p_2, p_2_pw4 = p_2_pw3 * p_2,
               d_12_pw2 = d_12 * d_12, b_pw2 = b * b;
  double fac[5];
  fac[0] = -f_2_pw2 * p_2_pw4 - p_2_pw4 * f_1_pw2 - p_2_pw4;
  fac[1] = 2 * p_2_pw3 * d_12 * b + 2 * f_2_pw2 * p_2_pw3 * d_12 * b - 2 * f_2 * p_2_pw3 * f_1 * d_12;
  fac[2] = -f_2_pw2 * p_2_pw2 * p_1_pw2 - f_2_pw2 * p_2_pw2 * d_12_pw2 * b_pw2 - f_2_pw2 * p_2_pw2 * d_12_pw2 +
           f_2_pw2 * p_2_pw4 + p_2_pw4 * f_1_pw2 + 2 * p_1 * p_2_pw2 * d_12 +
           2 * f_1 * f_2 * p_1 * p_2_pw2 * d_12 * b - p_2_pw2 * p_1_pw2 * f_1_pw2 +
           2 * p_1 * p_2_pw2 * f_2_pw2 * d_12 - p_2_pw2 * d_12_pw2 * b_pw2 - 2 * p_1_pw2 * p_2_pw2;
  fac[3] = 2 * p_1_pw2 * p_2 * d_12 * b + 2 * f_2 * p_2_pw3 * f_1 * d_12 - 2 * f_2_pw2 * p_2_pw3 * d_12 * b -
           2 * p_1 * p_2 * d_12_pw2 * b;
  fac[4] = -2 * f_2 * p_2_pw2 * f_1 * p_1 * d_12 * b + f_2_pw2 * p_2_pw2 * d_12_pw2 + 2 * p_1_pw3 * d_12 -
           p_1_pw2 * d_12_pw2 + f_2_pw2 * p_2_pw2 * p_1_pw2 - p_1_pw4 - 2 * f_2_pw2 * p_2_pw2 * p_1 * d_12 +
           p_2_pw2 * f_1_pw2 * p_1_pw2 + f_2_pw2 * p_2_pw2 * d_12_pw2 * b_pw2;
  pf_p3p_quartic(fac, F.roots);
  F.e1 = e1; F.e2 = e2; F.e3 = e3; F.n1 = n1; F.n2 = n2; F.n3 = n3; F.P1 = P1;
  F.f_1 = f_1; F.f_2 = f_2; F.p_1 = p_1; F.p_2 = p_2; F.d_12 = d_12; F.b = b;
  return true;
}

// Back-substitution of root i (p3p.cpp:181-233): S = [R | C] column-major 3x4, R = N^T R'^T T, C = P1 + N^T C'
__device__ __noinline__ void pf_p3p_backsub(const pf_p3p_frame& F, int i, double S[12])
{
  const double r = F.roots[i];
  const double cot_alpha = (-F.f_1 * F.p_1 / F.f_2 - r * F.p_2 + F.d_12 * F.b) / (-F.f_1 * r * F.p_2 / F.f_2 + F.p_1 - F.d_12);
  const double cos_theta = r, sin_theta = sqrt(1 - r * r);
  const double sin_alpha = sqrt(1 / (cot_alpha * cot_alpha + 1));
  double cos_alpha = sqrt(1 - sin_alpha * sin_alpha);
  if (cot_alpha < 0) cos_alpha = -cos_alpha;
  const double k = sin_alpha * F.b + cos_alpha;
  const double Cx = F.d_12 * cos_alpha * k, Cy = cos_theta * F.d_12 * sin_alpha * k, Cz = sin_theta * F.d_12 * sin_alpha * k;
  S[9] = F.P1.x + (F.n1.x * Cx + F.n2.x * Cy + F.n3.x * Cz);
  S[10] = F.P1.y + (F.n1.y * Cx + F.n2.y * Cy + F.n3.y * Cz);
  S[11] = F.P1.z + (F.n1.z * Cx + F.n2.z * Cy + F.n3.z * Cz);
  // R' as written at :205-214; M1 = N^T R'^T : M1(r,c) = sum_k N(k,r) R'(c,k)
  const double Rp[3][3] = {{-cos_alpha, -sin_alpha * cos_theta, -sin_alpha * sin_theta},
                           {sin_alpha, -cos_alpha * cos_theta, -cos_alpha * sin_theta},
                           {0.0, -sin_theta, cos_theta}};
  const double Nm[3][3] = {{F.n1.x, F.n1.y, F.n1.z}, {F.n2.x, F.n2.y, F.n2.z}, {F.n3.x, F.n3.y, F.n3.z}};
  const double Tm[3][3] = {{F.e1.x, F.e1.y, F.e1.z}, {F.e2.x, F.e2.y, F.e2.z}, {F.e3.x, F.e3.y, F.e3.z}};
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
  pf_p3p_frame F;
  const bool ok = pf_p3p_setup(pf_d3m(f[0], f[1], f[2]), pf_d3m(f[3], f[4], f[5]), pf_d3m(f[6], f[7], f[8]),
                               pf_d3m(w[0], w[1], w[2]), pf_d3m(w[3], w[4], w[5]), pf_d3m(w[6], w[7], w[8]), F);
  rc[i] = ok ? 0 : -1;
  if (!ok) return;
  for (int k = 0; k < 4; ++k) {
    double S[12];
    pf_p3p_backsub(F, k, S);
    for (int e = 0; e < 12; ++e) sol[48 * (size_t)i + 12 * k + e] = S[e];
  }
}

// ---- prologue: bearing vectors, cleared outputs ------------------------------------------------------------------
__global__ void pf_p3p_prep_kernel(pf_p3p_params p)
{
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < p.D; i += gridDim.x * blockDim.x) {
    pf_d3 v = pf_d3m((p.img[2 * i] - p.K[2]) / p.K[0], (p.img[2 * i + 1] - p.K[5]) / p.K[4], 1.0);
    v = pf_d3_divs(v, pf_d3_norm(v));
    p.vec[3 * i] = v.x; p.vec[3 * i + 1] = v.y; p.vec[3 * i + 2] = v.z;
  }
  // detections sorted by x (rank sort; a NaN column sorts last and is never inside a window)
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < p.D; i += gridDim.x * blockDim.x) {
    const double xi = p.img[2 * i], ki = xi != xi ? INFINITY : xi;
    int rank = 0;
    for (int j = 0; j < p.D; ++j) {
      const double xj = p.img[2 * j], kj = xj != xj ? INFINITY : xj;
      rank += (kj < ki || (kj == ki && j < i)) ? 1 : 0;
    }
    p.simg[2 * rank] = xi; p.simg[2 * rank + 1] = p.img[2 * i + 1];
    p.sidx[rank] = i;
  }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < p.D * p.L; i += gridDim.x * blockDim.x) p.hist[i] = 0u;
  if (blockIdx.x == 0 && threadIdx.x < 4) p.stats[threadIdx.x] = 0ull;
}

__device__ __forceinline__ double pf_sqd2(double ax, double ay, double bx, double by)
{
  const double dx = ax - bx, dy = ay - by;
  return dx * dx + dy * dy;
}

// ---- the vote (pose_estimator.cpp:1540-1717) ---------------------------------------------------------------------
// grid: one CTA per pair (i0 < i1); dynamic shared memory: hist [D*L] u32 | sxy [D] double2 | sid [D] int | pass [D] u8
__global__ void __launch_bounds__(PF_P3P_BLOCK) pf_p3p_hist_kernel(pf_p3p_params p)
{
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int D = p.D, L = p.L;
  double2* sxy = reinterpret_cast<double2*>(smem_raw);
  unsigned* sh = reinterpret_cast<unsigned*>(sxy + D);
  int* sid = reinterpret_cast<int*>(sh + D * L);
  unsigned char* pass = reinterpret_cast<unsigned char*>(sid + D);
  const int tid = threadIdx.x;

  // pair index -> (i0, i1), i0 < i1 < D - 1 (a third detection must follow)
  int i0 = 0, rem = blockIdx.x;
  while (rem >= D - 2 - i0) { rem -= D - 2 - i0; ++i0; }
  const int i1 = i0 + 1 + rem;

  for (int i = tid; i < D; i += PF_P3P_BLOCK) {
    sxy[i] = make_double2(p.simg[2 * i], p.simg[2 * i + 1]);
    sid[i] = p.sidx[i];
  }
  for (int i = tid; i < D * L; i += PF_P3P_BLOCK) sh[i] = 0u;
  __syncthreads();

  const double thresh = 10000.0 * 100.0;   // threshDist, threshDist2 (:1558-1559)
  const double d1x = p.img[2 * i0], d1y = p.img[2 * i0 + 1], d2x = p.img[2 * i1], d2y = p.img[2 * i1 + 1];
  unsigned n_triples = 0, n_solved = 0, n_proj = 0, n_vote = 0;

  // pre-filters of the triple (:1561-1581), once per i2
  for (int k = i1 + 1 + tid; k < D; k += PF_P3P_BLOCK) {
    const double d3x = p.img[2 * k], d3y = p.img[2 * k + 1];
    bool ok = !(pf_sqd2(d1x, d1y, d2x, d2y) > thresh) && !(pf_sqd2(d1x, d1y, d3x, d3y) > thresh) &&
              !(pf_sqd2(d2x, d2y, d3x, d3y) > thresh);
    if (ok) {
      const double mx = (d1x + d2x + d3x) / 3, my = (d1y + d2y + d3y) / 3;
      int cd = 0;
      for (int q = 0; q < D; ++q) cd += pf_sqd2(mx, my, sxy[q].x, sxy[q].y) < thresh ? 1 : 0;
      ok = cd >= 5;
    }
    pass[k] = ok ? 1 : 0;
    n_triples += ok ? 1u : 0u;
  }
  __syncthreads();

  const int nperm = L * (L - 1) * (L - 2);
  const int items = (D - 1 - i1) * nperm;
  const pf_d3 f1 = pf_d3m(p.vec[3 * i0], p.vec[3 * i0 + 1], p.vec[3 * i0 + 2]);
  const pf_d3 f2 = pf_d3m(p.vec[3 * i1], p.vec[3 * i1 + 1], p.vec[3 * i1 + 2]);
  const double tol = p.tol;
  const double reach = tol + 1e-6;   // column window; every candidate is then tested exactly
  const int n_un = L - 3;

  // The loop is written so that a warp stays converged: every thread runs the same number of steps, dead
  // threads (past the end, filtered triple, colinear LEDs) skip the bodies, and the warp re-joins after every
  // solution -- left to itself the compiler lets the threads that voted drift apart for the rest of the CTA.
  for (int base = 0; base < items; base += PF_P3P_BLOCK) {
    const int it = base + tid;
    const bool in_range = it < items;
    const int i2 = i1 + 1 + (in_range ? it : 0) / nperm;
    int pr = (in_range ? it : 0) % nperm;
    const int j0 = pr / ((L - 1) * (L - 2));
    pr -= j0 * (L - 1) * (L - 2);
    int j1 = pr / (L - 2), j2 = pr % (L - 2);
    j1 += j1 >= j0 ? 1 : 0;
    const int lo = j0 < j1 ? j0 : j1, hi = j0 < j1 ? j1 : j0;
    j2 += j2 >= lo ? 1 : 0;
    j2 += j2 >= hi ? 1 : 0;

    const pf_d3 f3 = pf_d3m(p.vec[3 * i2], p.vec[3 * i2 + 1], p.vec[3 * i2 + 2]);
    pf_p3p_frame F;
    bool live = in_range && pass[i2] != 0;
    if (live)
      live = pf_p3p_setup(f1, f2, f3, pf_d3m(p.obj[3 * j0], p.obj[3 * j0 + 1], p.obj[3 * j0 + 2]),
                          pf_d3m(p.obj[3 * j1], p.obj[3 * j1 + 1], p.obj[3 * j1 + 2]),
                          pf_d3m(p.obj[3 * j2], p.obj[3 * j2 + 1], p.obj[3 * j2 + 2]), F);
    __syncwarp();
    n_solved += live ? 1u : 0u;
    const double d3x = p.img[2 * i2], d3y = p.img[2 * i2 + 1];
    const double mx = (d1x + d2x + d3x) / 3, my = (d1y + d2y + d3y) / 3;

    double Sprev[12];
#pragma unroll 1
    for (int k = 0; k < 4; ++k) {
      double S[12];
      bool use = false;
      if (live) {
        pf_p3p_backsub(F, k, S);
        bool skip = false;
        if (k > 0) {   // :1659-1663: a solution is dropped when ANY entry equals the previous solution's
#pragma unroll
          for (int e = 0; e < 12; ++e) skip |= (S[e] - Sprev[e] == 0.0);
        }
        bool finite = true;   // isFinite(H) (:1670)
#pragma unroll
        for (int e = 0; e < 12; ++e) { finite &= ((S[e] - S[e]) == (S[e] - S[e])); Sprev[e] = S[e]; }
        use = !skip && finite;
      }
      __syncwarp();
      if (use) {
      ++n_proj;
      // inverse of [R C; 0 1] by cofactors: Ai = adj(R) / det(R), ti = -Ai C
      const double a00 = S[0], a10 = S[1], a20 = S[2], a01 = S[3], a11 = S[4], a21 = S[5], a02 = S[6], a12 = S[7], a22 = S[8];
      const double c00 = a11 * a22 - a12 * a21, c01 = a12 * a20 - a10 * a22, c02 = a10 * a21 - a11 * a20;
      const double det = a00 * c00 + a01 * c01 + a02 * c02;
      const double i00 = c00 / det, i01 = (a02 * a21 - a01 * a22) / det, i02 = (a01 * a12 - a02 * a11) / det;
      const double i10 = c01 / det, i11 = (a00 * a22 - a02 * a20) / det, i12 = (a02 * a10 - a00 * a12) / det;
      const double i20 = c02 / det, i21 = (a01 * a20 - a00 * a21) / det, i22 = (a00 * a11 - a01 * a10) / det;
      const double t0 = -(i00 * S[9] + i01 * S[10] + i02 * S[11]), t1 = -(i10 * S[9] + i11 * S[10] + i12 * S[11]),
                   t2 = -(i20 * S[9] + i21 * S[10] + i22 * S[11]);
      // project the unused LEDs (ascending LED number, :1637-1656) with the full K (project2d :1017-1034)
      double pu[PF_P3P_MAX_L - 3], pv[PF_P3P_MAX_L - 3];
      int pl[PF_P3P_MAX_L - 3];
      int m = 0;
      for (int ll = 0; ll < L; ++ll) {
        if (ll == j0 || ll == j1 || ll == j2) continue;
        const double X = p.obj[3 * ll], Y = p.obj[3 * ll + 1], Z = p.obj[3 * ll + 2];
        const double xc = i00 * X + i01 * Y + i02 * Z + t0, yc = i10 * X + i11 * Y + i12 * Z + t1,
                     zc = i20 * X + i21 * Y + i22 * Z + t2;
        const double q0 = p.K[0] * xc + p.K[1] * yc + p.K[2] * zc, q1 = p.K[3] * xc + p.K[4] * yc + p.K[5] * zc,
                     q2 = p.K[6] * xc + p.K[7] * yc + p.K[8] * zc;
        pu[m] = q0 / q2; pv[m] = q1 / q2; pl[m] = ll;
        ++m;
      }
      // votes (:1680-1712): every unused detection close to the triple's centroid whose nearest projected LED is
      // closer than tol votes for (detection, that LED); if any did, the triple's own three pairs vote too
      unsigned votes = 0;
      for (int a = 0; a < n_un; ++a) {
        const double u = pu[a], xlo = u - reach, xhi = u + reach;
        int lb = 0, ub = D;
        while (lb < ub) {   // first sorted detection with x >= xlo
          const int mid = (lb + ub) >> 1;
          if (sxy[mid].x < xlo) lb = mid + 1; else ub = mid;
        }
        for (int q = lb; q < D && sxy[q].x <= xhi; ++q) {
          const int det_i = sid[q];
          if (det_i == i0 || det_i == i1 || det_i == i2) continue;
          const double bx = sxy[q].x, by = sxy[q].y;
          if (!(pf_sqd2(mx, my, bx, by) < thresh)) continue;   // :1598
          double best = INFINITY;
          int bj = -1;
          for (int c = 0; c < n_un; ++c) {   // calculateMinDistancesAndPairs (:2093-2137): first strict minimum
            const double d2 = pf_sqd2(bx, by, pu[c], pv[c]);
            if (d2 < best) { best = d2; bj = c; }
          }
          if (bj == a && sqrt(best) < tol) {
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
      }   // use
      __syncwarp();
    }
  }
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
  return (size_t)D * 16 + (size_t)D * L * 4 + (size_t)D * 4 + (size_t)D + 16;
}

cudaError_t pf_p3p_launch_hist(const pf_p3p_params& p, cudaStream_t stream)
{
  const size_t smem = pf_p3p_smem_bytes(p.D, p.L);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(pf_p3p_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const unsigned pairs = (unsigned)((p.D - 1) * (p.D - 2) / 2);
  pf_p3p_prep_kernel<<<(p.D + 127) / 128, 128, 0, stream>>>(p);
  pf_p3p_hist_kernel<<<pairs, PF_P3P_BLOCK, smem, stream>>>(p);
  return cudaGetLastError();
}

cudaError_t pf_p3p_launch_solve(int n, const double* fv, const double* wp, double* sol, int* rc, cudaStream_t stream)
{
  pf_p3p_solve_kernel<<<(n + PF_P3P_BLOCK - 1) / PF_P3P_BLOCK, PF_P3P_BLOCK, 0, stream>>>(n, fv, wp, sol, rc);
  return cudaGetLastError();
}

