// pf_tail.cuh -- FP64 "tail" of a tracked frame: Gauss-Newton refinement of the winner
// particle, SE(3) exp/log maps and the constant-velocity prediction.  PF_HD: used by the
// host entry points pfmpe_optimise_pose / pfmpe_predict_pose (pfmpe_api.cu) and by the
// device-resident tracking kernel (pf_kernels.cuh).
//
// Reference semantics (file:line are /root/reference/pf_mpe_lib/src/pose_estimator.cpp):
//   optimisePose :1805-2009, computeJacobian :2163-2192, exponentialMap :2194-2226,
//   logarithmMap :2228-2296, predictPose :995-1010, updatePose :2011-2021.
// Poses here are affine 3x4 row-major double (bottom row 0 0 0 1 implied): a[4*i+j].
#pragma once
#include "pf_math.cuh"

struct pf_aff { double a[12]; };

PF_HD pf_aff pf_aff_identity()
{
  pf_aff T;
  PF_UNROLL
  for (int i = 0; i < 12; ++i) T.a[i] = 0.0;
  T.a[0] = T.a[5] = T.a[10] = 1.0;
  return T;
}

PF_HD pf_aff pf_aff_mul(const pf_aff& A, const pf_aff& B)
{
  pf_aff C;
  PF_UNROLL
  for (int i = 0; i < 3; ++i)
    PF_UNROLL
    for (int j = 0; j < 4; ++j) {
      double s = A.a[4 * i] * B.a[j] + A.a[4 * i + 1] * B.a[4 + j] + A.a[4 * i + 2] * B.a[8 + j];
      if (j == 3) s += A.a[4 * i + 3];
      C.a[4 * i + j] = s;
    }
  return C;
}

// true inverse of the affine map (the rotation block need not be exactly orthonormal)
PF_HD pf_aff pf_aff_inverse(const pf_aff& T)
{
  const double* a = T.a;
  double c00 = a[5] * a[10] - a[6] * a[9], c01 = a[6] * a[8] - a[4] * a[10], c02 = a[4] * a[9] - a[5] * a[8];
  double det = a[0] * c00 + a[1] * c01 + a[2] * c02;
  double id = 1.0 / det;
  pf_aff R;
  R.a[0] = c00 * id; R.a[1] = (a[2] * a[9] - a[1] * a[10]) * id; R.a[2] = (a[1] * a[6] - a[2] * a[5]) * id;
  R.a[4] = c01 * id; R.a[5] = (a[0] * a[10] - a[2] * a[8]) * id; R.a[6] = (a[2] * a[4] - a[0] * a[6]) * id;
  R.a[8] = c02 * id; R.a[9] = (a[1] * a[8] - a[0] * a[9]) * id;  R.a[10] = (a[0] * a[5] - a[1] * a[4]) * id;
  PF_UNROLL
  for (int i = 0; i < 3; ++i)
    R.a[4 * i + 3] = -(R.a[4 * i] * a[3] + R.a[4 * i + 1] * a[7] + R.a[4 * i + 2] * a[11]);
  return R;
}

// column-major 4x4 (Eigen layout at the ABI) <-> affine 3x4 row-major
PF_HD pf_aff pf_aff_from_cm16(const double* m)
{
  pf_aff T;
  PF_UNROLL
  for (int i = 0; i < 3; ++i)
    PF_UNROLL
    for (int j = 0; j < 4; ++j) T.a[4 * i + j] = m[i + 4 * j];
  return T;
}
PF_HD void pf_aff_to_cm16(const pf_aff& T, double* m)
{
  PF_UNROLL
  for (int i = 0; i < 3; ++i)
    PF_UNROLL
    for (int j = 0; j < 4; ++j) m[i + 4 * j] = T.a[4 * i + j];
  m[3] = 0.0; m[7] = 0.0; m[11] = 0.0; m[15] = 1.0;
}
PF_HD pf_pose pf_aff_to_f32(const pf_aff& T)
{
  pf_pose p;
  PF_UNROLL
  for (int i = 0; i < 12; ++i) p.p[i] = (float)T.a[i];
  return p;
}
PF_HD pf_aff pf_aff_from_f32(const pf_pose& p)
{
  pf_aff T;
  PF_UNROLL
  for (int i = 0; i < 12; ++i) T.a[i] = (double)p.p[i];
  return T;
}

// exponentialMap (:2194-2226): twist = (upsilon, omega)
PF_HD pf_aff pf_exp_map(const double* tw)
{
  const double wx = tw[3], wy = tw[4], wz = tw[5];
  const double theta = sqrt(wx * wx + wy * wy + wz * wz);
  const double th2 = theta * theta;
  // Omega and Omega^2 (row-major 3x3)
  const double O[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
  double O2[9];
  PF_UNROLL
  for (int i = 0; i < 3; ++i)
    PF_UNROLL
    for (int j = 0; j < 3; ++j) O2[3 * i + j] = O[3 * i] * O[j] + O[3 * i + 1] * O[3 + j] + O[3 * i + 2] * O[6 + j];
  double R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, V[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (theta != 0) {
    const double st = sin(theta), ct = cos(theta);
    PF_UNROLL
    for (int i = 0; i < 9; ++i) {
      R[i] += O[i] / theta * st + O2[i] / th2 * (1 - ct);
      V[i] += (1 - ct) / th2 * O[i] + (theta - st) / (th2 * theta) * O2[i];
    }
  }
  pf_aff T;
  PF_UNROLL
  for (int i = 0; i < 3; ++i) {
    PF_UNROLL
    for (int j = 0; j < 3; ++j) T.a[4 * i + j] = R[3 * i + j];
    T.a[4 * i + 3] = V[3 * i] * tw[0] + V[3 * i + 1] * tw[1] + V[3 * i + 2] * tw[2];
  }
  return T;
}

// exponentialMap for the Gauss-Newton update (:1876), where the twist is a correction step: the divisions of the reference
// expressions (O / theta * sin, O^2 / theta^2 * (1 - cos), ...) are hoisted into one reciprocal, and below |omega| = 0.01 the
// three coefficients come from their Taylor series in theta^2 (next omitted term < 1e-21 relative; also free of the
// cancellation in (1 - cos theta) / theta^2).  Agrees with pf_exp_map to a few ulp; the dependent chain is several times shorter.
PF_HD pf_aff pf_exp_map_step(const double* tw)
{
  const double wx = tw[3], wy = tw[4], wz = tw[5];
  const double th2 = wx * wx + wy * wy + wz * wz;
  double a, b, c;                                  // R = I + a O + b O^2 ; V = I + b O + c O^2
  if (th2 < 1e-4) {
    a = 1.0 + th2 * (-1.0 / 6.0 + th2 * (1.0 / 120.0 + th2 * (-1.0 / 5040.0 + th2 * (1.0 / 362880.0))));
    b = 0.5 + th2 * (-1.0 / 24.0 + th2 * (1.0 / 720.0 + th2 * (-1.0 / 40320.0 + th2 * (1.0 / 3628800.0))));
    c = 1.0 / 6.0 + th2 * (-1.0 / 120.0 + th2 * (1.0 / 5040.0 + th2 * (-1.0 / 362880.0 + th2 * (1.0 / 39916800.0))));
  } else {
    const double theta = sqrt(th2);
    const double st = sin(theta), ct = cos(theta);
    const double it = 1.0 / theta, it2 = it * it;
    a = st * it; b = (1.0 - ct) * it2; c = (theta - st) * it2 * it;
  }
  const double O[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
  double O2[9];
  PF_UNROLL
  for (int i = 0; i < 3; ++i)
    PF_UNROLL
    for (int j = 0; j < 3; ++j) O2[3 * i + j] = O[3 * i] * O[j] + O[3 * i + 1] * O[3 + j] + O[3 * i + 2] * O[6 + j];
  pf_aff T;
  PF_UNROLL
  for (int i = 0; i < 3; ++i) {
    double V[3];
    PF_UNROLL
    for (int j = 0; j < 3; ++j) {
      const double id = i == j ? 1.0 : 0.0;
      T.a[4 * i + j] = id + a * O[3 * i + j] + b * O2[3 * i + j];
      V[j] = id + b * O[3 * i + j] + c * O2[3 * i + j];
    }
    T.a[4 * i + 3] = V[0] * tw[0] + V[1] * tw[1] + V[2] * tw[2];
  }
  return T;
}

// logarithmMap (:2228-2296), including Eigen's isApprox / isApproxToConstant tests
PF_HD void pf_log_map(const pf_aff& T, double* xi)
{
  const double* a = T.a;
  double dev = 0, nrm = 0;
  PF_UNROLL
  for (int i = 0; i < 3; ++i)
    PF_UNROLL
    for (int j = 0; j < 3; ++j) {
      double e = a[4 * i + j] - (i == j ? 1.0 : 0.0);
      dev += e * e;
      nrm += a[4 * i + j] * a[4 * i + j];
    }
  double wh[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  const bool near_identity = dev <= 1e-20 * (nrm < 3.0 ? nrm : 3.0);
  if (!near_identity) {
    double c = (a[0] + a[5] + a[10] - 1) / 2;
    c = c > 1 ? 1 : (c < -1 ? -1 : c);
    const double phi = acos(c);
    if (phi != 0) {
      const double f = 2 * sin(phi);
      PF_UNROLL
      for (int i = 0; i < 3; ++i)
        PF_UNROLL
        for (int j = 0; j < 3; ++j) wh[3 * i + j] = (a[4 * i + j] - a[4 * j + i]) / f * phi;
    }
  }
  const double w[3] = {wh[7], wh[2], wh[3]};   // (w_hat(2,1), w_hat(0,2), w_hat(1,0))
  const double wn = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  const double t[3] = {a[3], a[7], a[11]};
  double Ai[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  const bool t_zero = (t[0] * t[0] + t[1] * t[1] + t[2] * t[2]) <= 0.0;  // isApproxToConstant(0): exact zero only
  if (!t_zero) {
    Ai[0] = Ai[4] = Ai[8] = 1.0;
    if (!(wn == 0 || sin(wn) == 0)) {
      const double f = (2 * sin(wn) - wn * (1 + cos(wn))) / (2 * wn * wn * sin(wn));
      PF_UNROLL
      for (int i = 0; i < 3; ++i)
        PF_UNROLL
        for (int j = 0; j < 3; ++j) {
          double w2 = wh[3 * i] * wh[j] + wh[3 * i + 1] * wh[3 + j] + wh[3 * i + 2] * wh[6 + j];
          Ai[3 * i + j] += -wh[3 * i + j] / 2 + f * w2;
        }
    }
  }
  PF_UNROLL
  for (int i = 0; i < 3; ++i) xi[i] = Ai[3 * i] * t[0] + Ai[3 * i + 1] * t[1] + Ai[3 * i + 2] * t[2];
  xi[3] = w[0]; xi[4] = w[1]; xi[5] = w[2];
}

// predictPose (:995-1010): P = exp( log(prev^-1 cur) / (t_cur - t_prev) * (t_pred - t_cur) ), pred = cur P
PF_HD void pf_predict_pose(const pf_aff& prev, const pf_aff& cur, double t_prev, double t_cur, double t_pred,
                           pf_aff& P, pf_aff& pred)
{
  double d[6];
  pf_log_map(pf_aff_mul(pf_aff_inverse(prev), cur), d);
  PF_UNROLL
  for (int i = 0; i < 6; ++i) d[i] = d[i] / (t_cur - t_prev) * (t_pred - t_cur);
  P = pf_exp_map(d);
  pred = pf_aff_mul(cur, P);
}

// 6x6 symmetric solve, LDL^T with diagonal pivoting (Eigen's A.ldlt().solve(b), :1872).
// A: full 6x6 (symmetric), row-major.
PF_HD void pf_ldlt6_solve(const double* A, const double* b, double* x)
{
  double M[36];
  int p[6];
  PF_UNROLL
  for (int i = 0; i < 36; ++i) M[i] = A[i];
  PF_UNROLL
  for (int i = 0; i < 6; ++i) p[i] = i;
  PF_UNROLL
  for (int k = 0; k < 6; ++k) {
    int piv = k;
    double big = fabs(M[7 * k]);
    PF_UNROLL
    for (int i = k + 1; i < 6; ++i)
      if (fabs(M[7 * i]) > big) { big = fabs(M[7 * i]); piv = i; }
    if (piv != k) {
      PF_UNROLL
      for (int j = 0; j < 6; ++j) { double t = M[6 * k + j]; M[6 * k + j] = M[6 * piv + j]; M[6 * piv + j] = t; }
      PF_UNROLL
      for (int j = 0; j < 6; ++j) { double t = M[6 * j + k]; M[6 * j + k] = M[6 * j + piv]; M[6 * j + piv] = t; }
      int t = p[k]; p[k] = p[piv]; p[piv] = t;
    }
    const double d = M[7 * k];
    if (d == 0.0) continue;
    PF_UNROLL
    for (int i = k + 1; i < 6; ++i) M[6 * i + k] /= d;       // L(i,k)
    PF_UNROLL
    for (int i = k + 1; i < 6; ++i)
      PF_UNROLL
      for (int j = k + 1; j <= i; ++j) {
        M[6 * i + j] -= M[6 * i + k] * d * M[6 * j + k];
        M[6 * j + i] = M[6 * i + j];
      }
  }
  double y[6];
  PF_UNROLL
  for (int i = 0; i < 6; ++i) {
    double s = b[p[i]];
    PF_UNROLL
    for (int k = 0; k < i; ++k) s -= M[6 * i + k] * y[k];
    y[i] = s;
  }
  PF_UNROLL
  for (int i = 0; i < 6; ++i) y[i] = M[7 * i] != 0.0 ? y[i] / M[7 * i] : 0.0;
  PF_UNROLL
  for (int i = 5; i >= 0; --i) {
    double s = y[i];
    PF_UNROLL
    for (int k = i + 1; k < 6; ++k) s -= M[6 * k + i] * y[k];
    y[i] = s;
  }
  PF_UNROLL
  for (int i = 0; i < 6; ++i) x[p[i]] = y[i];
}

// 6x6 inverse by Gauss-Jordan with partial pivoting (pose_covariance_ = A.inverse(), :2004)
PF_HD void pf_inv6(const double* A, double* Ai)
{
  double M[36];
  PF_UNROLL
  for (int i = 0; i < 36; ++i) { M[i] = A[i]; Ai[i] = 0.0; }
  PF_UNROLL
  for (int i = 0; i < 6; ++i) Ai[7 * i] = 1.0;
  PF_UNROLL
  for (int c = 0; c < 6; ++c) {
    int piv = c;
    double big = fabs(M[6 * c + c]);
    PF_UNROLL
    for (int r = c + 1; r < 6; ++r)
      if (fabs(M[6 * r + c]) > big) { big = fabs(M[6 * r + c]); piv = r; }
    if (piv != c) {
      PF_UNROLL
      for (int j = 0; j < 6; ++j) {
        double t = M[6 * c + j]; M[6 * c + j] = M[6 * piv + j]; M[6 * piv + j] = t;
        t = Ai[6 * c + j]; Ai[6 * c + j] = Ai[6 * piv + j]; Ai[6 * piv + j] = t;
      }
    }
    const double s = 1.0 / M[6 * c + c];
    PF_UNROLL
    for (int j = 0; j < 6; ++j) { M[6 * c + j] *= s; Ai[6 * c + j] *= s; }
    PF_UNROLL
    for (int r = 0; r < 6; ++r) {
      if (r == c) continue;
      const double f = M[6 * r + c];
      if (f == 0.0) continue;
      PF_UNROLL
      for (int j = 0; j < 6; ++j) { M[6 * r + j] -= f * M[6 * c + j]; Ai[6 * r + j] -= f * Ai[6 * c + j]; }
    }
  }
}

// ---- 6x6 SPD system in 3x3 blocks ------------------------------------------------------------
// A = [P Q; Q^T R] (symmetric positive definite normal matrix of the Gauss-Newton step).
// inv(P) and inv(S), S = R - Q^T inv(P) Q, by cofactors: two reciprocals on the critical path
// instead of six pivots (the reference solves with Eigen's pivoted LDL^T, :1872; same solution up
// to rounding).  Device: evaluated redundantly by every lane from the shared-memory copy of A; host: pf_optimise_pose.
// pf_blk6_factor returns false when a block is singular (the host form then takes the pivoted LDL^T for that step).
struct pf_blk6 { double Pi[6], W[9], Si[6]; };     // symmetric 3x3 stored as 00 01 02 11 12 22

PF_HD bool pf_sym3_inv(const double* m, double* o)
{
  const double c00 = m[3] * m[5] - m[4] * m[4], c01 = m[2] * m[4] - m[1] * m[5], c02 = m[1] * m[4] - m[2] * m[3];
  const double id = 1.0 / (m[0] * c00 + m[1] * c01 + m[2] * c02);
  o[0] = c00 * id; o[1] = c01 * id; o[2] = c02 * id;
  o[3] = (m[0] * m[5] - m[2] * m[2]) * id; o[4] = (m[1] * m[2] - m[0] * m[4]) * id;
  o[5] = (m[0] * m[3] - m[1] * m[1]) * id;
  return id - id == 0.0;                                 // false: singular block (1 / 0) or NaN
}
#define PF_S3(m, i, j) ((m)[(i) <= (j) ? ((i) == 0 ? (j) : ((i) == 1 ? 2 + (j) : 5)) : ((j) == 0 ? (i) : ((j) == 1 ? 2 + (i) : 5))])

PF_HD bool pf_blk6_factor(const double* A, pf_blk6& f)
{
  const double P[6] = {A[0], A[1], A[2], A[7], A[8], A[14]};
  const bool okP = pf_sym3_inv(P, f.Pi);
PF_UNROLL
  for (int i = 0; i < 3; ++i)
PF_UNROLL
    for (int j = 0; j < 3; ++j)   // W = inv(P) Q,  Q(k,j) = A[6k + 3 + j]
      f.W[3 * i + j] = PF_S3(f.Pi, i, 0) * A[3 + j] + PF_S3(f.Pi, i, 1) * A[9 + j] + PF_S3(f.Pi, i, 2) * A[15 + j];
  double S[6];
  int t = 0;
PF_UNROLL
  for (int i = 0; i < 3; ++i)
PF_UNROLL
    for (int j = i; j < 3; ++j, ++t)   // S = R - Q^T W
      S[t] = A[6 * (3 + i) + 3 + j] - (A[3 + i] * f.W[j] + A[9 + i] * f.W[3 + j] + A[15 + i] * f.W[6 + j]);
  return pf_sym3_inv(S, f.Si) && okP;
}

PF_HD void pf_blk6_solve(const pf_blk6& f, const double* A, const double* b, double* x)
{
  double y1[3], r2[3];
PF_UNROLL
  for (int i = 0; i < 3; ++i) y1[i] = PF_S3(f.Pi, i, 0) * b[0] + PF_S3(f.Pi, i, 1) * b[1] + PF_S3(f.Pi, i, 2) * b[2];
PF_UNROLL
  for (int i = 0; i < 3; ++i) r2[i] = b[3 + i] - (A[3 + i] * y1[0] + A[9 + i] * y1[1] + A[15 + i] * y1[2]);
PF_UNROLL
  for (int i = 0; i < 3; ++i) x[3 + i] = PF_S3(f.Si, i, 0) * r2[0] + PF_S3(f.Si, i, 1) * r2[1] + PF_S3(f.Si, i, 2) * r2[2];
PF_UNROLL
  for (int i = 0; i < 3; ++i) x[i] = y1[i] - (f.W[3 * i] * x[3] + f.W[3 * i + 1] * x[4] + f.W[3 * i + 2] * x[5]);
}

// inv(A) = [inv(P) + W Si W^T, -W Si; -Si W^T, Si]  -> cov (row-major 6x6)
PF_HD void pf_blk6_inverse(const pf_blk6& f, double* cov)
{
  double WS[9];
PF_UNROLL
  for (int i = 0; i < 3; ++i)
PF_UNROLL
    for (int j = 0; j < 3; ++j)
      WS[3 * i + j] = f.W[3 * i] * PF_S3(f.Si, 0, j) + f.W[3 * i + 1] * PF_S3(f.Si, 1, j) + f.W[3 * i + 2] * PF_S3(f.Si, 2, j);
PF_UNROLL
  for (int i = 0; i < 3; ++i)
PF_UNROLL
    for (int j = 0; j < 3; ++j) {
      cov[6 * i + j] = PF_S3(f.Pi, i, j) + (WS[3 * i] * f.W[3 * j] + WS[3 * i + 1] * f.W[3 * j + 1] + WS[3 * i + 2] * f.W[3 * j + 2]);
      cov[6 * i + 3 + j] = -WS[3 * i + j];
      cov[6 * (3 + j) + i] = -WS[3 * i + j];
      cov[6 * (3 + i) + 3 + j] = PF_S3(f.Si, i, j);
    }
}

// optimisePose (:1805-2009).  K9 row-major; obj: L x 3 doubles; img: D x 2 doubles;
// pairs: npairs x (LED+1, blob+1).  pose in/out.  cov: 6x6 (symmetric; row-major == col-major
// up to rounding).  Returns PubData.numIter (-1 when max_itr is hit).
PF_HD int pf_optimise_pose(const double* K9, const double* obj, const double* img, const unsigned* pairs,
                           int npairs, pf_aff& pose, double* cov)
{
  const double conv = 1e-13;
  const int max_itr = 500;
  const pf_aff pose0 = pose;
  const double fx = K9[0], fy = K9[4];
  double e_first = 0, e_last = 0;
  double A[36], b[6], dT[6];
  PF_UNROLL
  for (int i = 0; i < 36; ++i) A[i] = 0.0;
  pf_blk6 f;
  bool blocks = false;
  int reported = -1;
  for (int it = 0; it < max_itr; ++it) {
    PF_UNROLL
    for (int i = 0; i < 36; ++i) A[i] = 0.0;
    PF_UNROLL
    for (int i = 0; i < 6; ++i) b[i] = 0.0;
    for (int q = 0; q < npairs; ++q) {
      const unsigned led = pairs[2 * q], det = pairs[2 * q + 1];
      if (det == 0) continue;   // :1847
      const double* X = obj + 3 * (led - 1);
      const double* a = pose.a;
      const double x = a[0] * X[0] + a[1] * X[1] + a[2] * X[2] + a[3];
      const double y = a[4] * X[0] + a[5] * X[1] + a[6] * X[2] + a[7];
      const double z = a[8] * X[0] + a[9] * X[1] + a[10] * X[2] + a[11];
      // project2d with the full K (:1017-1034)
      const double p0 = K9[0] * x + K9[1] * y + K9[2] * z;
      const double p1 = K9[3] * x + K9[4] * y + K9[5] * z;
      const double p2 = K9[6] * x + K9[7] * y + K9[8] * z;
      const double ip2 = 1.0 / p2, iz = 1.0 / z;
      const double ex = img[2 * (det - 1)] - p0 * ip2, ey = img[2 * (det - 1) + 1] - p1 * ip2;
      if (it == 0) e_first = sqrt(ex * ex + ey * ey);                 // `=+` in the reference: last pair wins (:1859)
      else if (it + 1 == max_itr) e_last = sqrt(ex * ex + ey * ey);   // :1861
      // computeJacobian (:2163-2192) with 1/z formed once (as the device form, pf_gn_warp)
      const double xz = x * iz, yz = y * iz;
      const double J0[6] = {fx * iz, 0, -fx * xz * iz, -fx * xz * yz, fx * (1.0 + xz * xz), -fx * yz};
      const double J1[6] = {0, fy * iz, -fy * yz * iz, -fy * (1.0 + yz * yz), fy * xz * yz, fy * xz};
      PF_UNROLL
      for (int r = 0; r < 6; ++r) {
        PF_UNROLL
        for (int c = r; c < 6; ++c) A[6 * r + c] += J0[r] * J0[c] + J1[r] * J1[c];   // upper triangle, mirrored below
        b[r] += J0[r] * ex + J1[r] * ey;
      }
    }
    PF_UNROLL
    for (int r = 1; r < 6; ++r) {
      PF_UNROLL
      for (int c = 0; c < r; ++c) A[6 * r + c] = A[6 * c + r];
    }
    // A.ldlt().solve(b) (:1872): in 3x3 blocks like the device form (a third of the time of the pivoted LDL^T, the same
    // solution up to rounding); a singular block (degenerate correspondences) takes the pivoted LDL^T, which skips zero pivots
    blocks = pf_blk6_factor(A, f);
    if (blocks) pf_blk6_solve(f, A, b, dT); else pf_ldlt6_solve(A, b, dT);
    pose = pf_aff_mul(pf_exp_map_step(dT), pose);   // :1876
    double nm = -1;
    PF_UNROLL
    for (int i = 0; i < 6; ++i) { double v = fabs(dT[i]); if (v > nm) nm = v; }
    if (nm <= conv) { reported = it; break; }
    if (it + 1 == max_itr && e_first < e_last) pose = pose0;   // :1886-1892
  }
  if (blocks) pf_blk6_inverse(f, cov); else pf_inv6(A, cov);   // pose_covariance_ = A.inverse() (:2004)
  return reported;
}

// Per-tracker constants of the frame about to be processed.  Written by the host for
// pfmpe_update (per-frame mode) or by the device tail of the previous frame (track mode).
struct pf_frame_const {
  float Tcur[12], Tpred[12], P[12], C[12];   // affine 3x4 row-major (current_pose_, predicted_pose_, predictionMatrix, camMoveInv)
  pf_noise_cfg noise;                        // ranges already multiplied by fac* (:488-516)
  uint32_t seed_state;                       // minstd_rand0 state after seeding (:476-477, injected seed)
  uint32_t isi;                              // it_since_initialized_
  int32_t D;                                 // numLED = image_points_.size()
  uint32_t flags;                            // PF_FC_*
};
#define PF_FC_CAM_IDENTITY 1u                /* camMoveInv is exactly I: C*E == E bit for bit, product skipped */
#define PF_FC_SMALL_ANG 2u                   /* every angle draw of the frame is within +-0.3 rad              */
#define PF_FC_TINY_ANG 4u                    /* ... within +-0.03 rad (the default noise ranges)               */

// ---------------------------------------------------------------------------------------
// frame constants from tracker inputs (host: pfmpe_update; device: track-mode tail)
// :488-516 -- fac* and the six uniform ranges (Q1: y,z reuse predictionMatrix(0,3))
// ---------------------------------------------------------------------------------------
PF_HD void pf_make_frame_const(pf_frame_const& fc, const pf_aff& cur, const pf_aff& pred, const pf_aff& P,
                               const pf_aff& C, uint32_t isi, double dt, uint32_t seed, int D,
                               double min_ang, double max_ang, double min_tr, double max_tr)
{
  PF_UNROLL
  for (int i = 0; i < 12; ++i) {
    fc.Tcur[i] = (float)cur.a[i];
    fc.Tpred[i] = (float)pred.a[i];
    fc.P[i] = (float)P.a[i];
    fc.C[i] = (float)C.a[i];
  }
  double fT, fR;
  if (isi == 1) { fT = 1.0; fR = 1.0; }
  else {
    double v = fabs(P.a[3]) / dt;
    v = v > 0.2 ? v : 0.2;        // std::max(0.2, .)
    v = v < 1.0 ? v : 1.0;        // std::min(., 1.0)
    fT = v / 4;
    fR = 0.2;
  }
  const double alo = min_ang * fR, ahi = max_ang * fR, tlo = min_tr * fT, thi = max_tr * fT;
  fc.noise.ang_lo = (float)alo; fc.noise.ang_span = (float)(ahi - alo);
  fc.noise.tr_lo = (float)tlo;  fc.noise.tr_span = (float)(thi - tlo);
  uint32_t x = seed % PF_RNG_M;
  fc.seed_state = x ? x : 1u;
  fc.isi = isi;
  fc.D = D;
  bool ident = true;
  for (int i = 0; i < 12; ++i) ident = ident && (C.a[i] == ((i == 0 || i == 5 || i == 10) ? 1.0 : 0.0));
  const double amax = (fabs(alo) > fabs(ahi) ? fabs(alo) : fabs(ahi)) * (1.0 + 0.025 * 7);   // scale <= 1.175 (:563)
  fc.flags = (ident ? PF_FC_CAM_IDENTITY : 0u) | (amax <= 0.3 ? PF_FC_SMALL_ANG : 0u) | (amax <= 0.03 ? PF_FC_TINY_ANG : 0u);
}

