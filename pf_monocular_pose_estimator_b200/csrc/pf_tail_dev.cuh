// pf_tail_dev.cuh -- warp-cooperative FP64 tail of a tracked frame (device only).
//
// Same semantics as the scalar PF_HD functions of pf_tail.cuh (which the host entry points
// pfmpe_optimise_pose / pfmpe_predict_pose use), restructured for one warp:
//   optimisePose :1805-2009  -> pf_gn_warp: lane q owns correspondence q (projection, residual,
//                               Jacobian rows); 27 lanes own the unique entries of A = sum J^T J and
//                               b = sum J^T e; the pivoted LDL^T (Eigen's A.ldlt().solve(b), :1872)
//                               runs on a 6x6 tile in shared memory with 21 lanes on the lower
//                               triangle; the triangular solves travel by warp shuffle.
//   A.inverse() :2004        -> pf_inv6_warp: Gauss-Jordan, lane c holds column c of [A | I].
// (file:line are /root/reference/pf_mpe_lib/src/pose_estimator.cpp)
// Everything that is O(1) and scalar (exp/log map, pose products) is evaluated redundantly by
// every lane, so the refined pose is available in registers of all lanes without broadcasts.
#pragma once
#include "pf_tail.cuh"

#if defined(__CUDACC__)

struct pf_tail_smem {
  double J[PF_MAX_LED][14];   // per correspondence: J row 0 (6), J row 1 (6), ex, ey
  double M[36];               // LDL^T work tile; afterwards the covariance (row-major)
  double A[36];               // sum J^T J of the last iteration
  double b[6];
  double dT[6];
  int perm[6];
  int pad[2];
};

#define PF_FULL 0xFFFFFFFFu

__device__ __forceinline__ double pf_shfl_d(double v, int src) { return __shfl_sync(PF_FULL, v, src); }

// lower-triangle lane map: lane l < 21 -> (i, j), j <= i
__device__ __forceinline__ void pf_tri_ij(int l, int& i, int& j)
{
  i = 0;
#pragma unroll
  for (int r = 1; r < 6; ++r) if (l >= r * (r + 1) / 2) i = r;
  j = l - i * (i + 1) / 2;
}

// Solve A x = b by LDL^T with diagonal pivoting; A (symmetric) in ts.M (full 6x6, overwritten),
// b in ts.b, x -> ts.dT.  All 32 lanes must call.
__device__ __forceinline__ void pf_ldlt6_warp(pf_tail_smem& ts, int lane)
{
  int ti, tj;
  pf_tri_ij(lane < 21 ? lane : 0, ti, tj);
  const bool tri = lane < 21;
  if (lane < 6) ts.perm[lane] = lane;
  __syncwarp();
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    int piv = k;
    double big = fabs(ts.M[7 * k]);
#pragma unroll
    for (int t = k + 1; t < 6; ++t) {
      const double v = fabs(ts.M[7 * t]);
      if (v > big) { big = v; piv = t; }
    }
    if (piv != k) {   // symmetric row/column swap k <-> piv (warp-uniform branch)
      const int pi = ti == k ? piv : (ti == piv ? k : ti);
      const int pj = tj == k ? piv : (tj == piv ? k : tj);
      const double v = tri ? ts.M[6 * pi + pj] : 0.0;
      __syncwarp();
      if (tri) { ts.M[6 * ti + tj] = v; ts.M[6 * tj + ti] = v; }
      if (lane == 0) { const int t = ts.perm[k]; ts.perm[k] = ts.perm[piv]; ts.perm[piv] = t; }
      __syncwarp();
    }
    const double d = ts.M[7 * k];
    if (d == 0.0) continue;
    if (tri && tj == k && ti > k) ts.M[6 * ti + k] = ts.M[6 * ti + k] / d;   // L(i,k)
    __syncwarp();
    if (tri && tj > k) {   // ti >= tj > k
      const double v = ts.M[6 * ti + tj] - ts.M[6 * ti + k] * d * ts.M[6 * tj + k];
      ts.M[6 * ti + tj] = v;
      ts.M[6 * tj + ti] = v;
    }
    __syncwarp();
  }
  // L y = P b ; D z = y ; L^T x' = z ; x = P^T x'
  const int li = lane < 6 ? lane : 0;
  double y = lane < 6 ? ts.b[ts.perm[li]] : 0.0;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    const double yk = pf_shfl_d(y, k);
    if (lane > k && lane < 6) y -= ts.M[6 * li + k] * yk;
  }
  if (lane < 6) { const double dd = ts.M[7 * li]; y = dd != 0.0 ? y / dd : 0.0; }
#pragma unroll
  for (int k = 5; k >= 1; --k) {
    const double yk = pf_shfl_d(y, k);
    if (lane < k) y -= ts.M[6 * k + li] * yk;
  }
  if (lane < 6) ts.dT[ts.perm[li]] = y;
  __syncwarp();
}

// cov = A^-1 (Gauss-Jordan with partial pivoting on [A | I]); A in ts.A, result -> ts.M row-major.
__device__ __forceinline__ void pf_inv6_warp(pf_tail_smem& ts, int lane)
{
  double col[6];
#pragma unroll
  for (int r = 0; r < 6; ++r) col[r] = lane < 6 ? ts.A[6 * r + lane] : ((lane - 6) == r ? 1.0 : 0.0);
#pragma unroll
  for (int c = 0; c < 6; ++c) {
    int piv = c;
    double big = fabs(col[c]);
#pragma unroll
    for (int r = c + 1; r < 6; ++r) {
      const double v = fabs(col[r]);
      if (v > big) { big = v; piv = r; }
    }
    piv = __shfl_sync(PF_FULL, piv, c);
#pragma unroll
    for (int r = c + 1; r < 6; ++r)
      if (piv == r) { const double t = col[c]; col[c] = col[r]; col[r] = t; }
    const double s = pf_shfl_d(1.0 / col[c], c);
    col[c] *= s;
#pragma unroll
    for (int r = 0; r < 6; ++r) {
      if (r == c) continue;
      const double f = pf_shfl_d(col[r], c);
      if (f != 0.0) col[r] -= f * col[c];
    }
  }
  __syncwarp();
  if (lane >= 6 && lane < 12) {
#pragma unroll
    for (int r = 0; r < 6; ++r) ts.M[6 * r + (lane - 6)] = col[r];
  }
  __syncwarp();
}

// optimisePose (:1805-2009) by one warp.  K9: row-major K (kernel parameter space); obj: L x 3
// marker coordinates (shared memory); img: detections of the frame (D x 2 doubles, global);
// pairs: (LED+1, blob+1) rows in registers.  pose in/out (replicated in every lane).  The
// covariance is left in ts.M (row-major).  Returns PubData.numIter (-1 when max_itr is hit).
template <int L>
__device__ __forceinline__ int pf_gn_warp(const double* K9, const double* obj, const double* img, const unsigned* pairs,
                                          int npairs, pf_aff& pose, pf_tail_smem& ts, int lane)
{
  const double conv = 1e-13;
  const int max_itr = 500;
  const pf_aff pose0 = pose;
  const double fx = K9[0], fy = K9[4];
  // my correspondence
  unsigned led = 0, det = 0;
#pragma unroll
  for (int q = 0; q < L; ++q)
    if (q == lane && q < npairs) { led = pairs[2 * q]; det = pairs[2 * q + 1]; }
  const bool active = lane < npairs && det != 0;   // :1847
  double X0 = 0, X1 = 0, X2 = 0, zx = 0, zy = 0;
  if (active) {
    X0 = obj[3 * (led - 1)]; X1 = obj[3 * (led - 1) + 1]; X2 = obj[3 * (led - 1) + 2];
    zx = img[2 * (det - 1)]; zy = img[2 * (det - 1) + 1];
  }
  const unsigned act_mask = __ballot_sync(PF_FULL, active);
  const int last_active = act_mask ? 31 - __clz(act_mask) : 0;
  // entry owned in the accumulation step: lanes 0..20 -> A(i,j), lanes 21..26 -> b(r)
  int ti, tj;
  pf_tri_ij(lane < 21 ? lane : 0, ti, tj);
  const bool own_A = lane < 21, own_b = lane >= 21 && lane < 27;
  const int rb = own_b ? lane - 21 : 0;
  const int i0 = own_A ? ti : rb, i1 = own_A ? tj : 12, i2 = own_A ? 6 + ti : 6 + rb, i3 = own_A ? 6 + tj : 13;

  double e_first = 0, e_last = 0;
  int reported = -1;
  for (int it = 0; it < max_itr; ++it) {
    double Jr[14];
#pragma unroll
    for (int i = 0; i < 14; ++i) Jr[i] = 0.0;
    double en = 0.0;
    if (active) {
      const double* a = pose.a;
      const double x = a[0] * X0 + a[1] * X1 + a[2] * X2 + a[3];
      const double y = a[4] * X0 + a[5] * X1 + a[6] * X2 + a[7];
      const double z = a[8] * X0 + a[9] * X1 + a[10] * X2 + a[11];
      const double p0 = K9[0] * x + K9[1] * y + K9[2] * z;     // project2d with the full K (:1017-1034)
      const double p1 = K9[3] * x + K9[4] * y + K9[5] * z;
      const double p2 = K9[6] * x + K9[7] * y + K9[8] * z;
      const double ex = zx - p0 / p2, ey = zy - p1 / p2;
      en = sqrt(ex * ex + ey * ey);
      const double z2 = z * z;
      Jr[0] = 1 / z * fx;  Jr[1] = 0;           Jr[2] = -x / z2 * fx;
      Jr[3] = -x * y / z2 * fx;  Jr[4] = (1 + (x * x / z2)) * fx;  Jr[5] = -y / z * fx;
      Jr[6] = 0;           Jr[7] = 1 / z * fy;  Jr[8] = -y / z2 * fy;
      Jr[9] = -(1 + y * y / z2) * fy;  Jr[10] = x * y / z2 * fy;   Jr[11] = x / z * fy;
      Jr[12] = ex; Jr[13] = ey;
    }
    if (it == 0) e_first = pf_shfl_d(en, last_active);                 // `=+`: last pair wins (:1859)
    else if (it + 1 == max_itr) e_last = pf_shfl_d(en, last_active);   // :1861
    if (lane < L) {
#pragma unroll
      for (int i = 0; i < 14; ++i) ts.J[lane][i] = Jr[i];
    }
    __syncwarp();
    if (lane < 27) {
      double acc = 0.0;
#pragma unroll
      for (int q = 0; q < L; ++q) acc += ts.J[q][i0] * ts.J[q][i1] + ts.J[q][i2] * ts.J[q][i3];
      if (own_A) { ts.A[6 * ti + tj] = acc; ts.A[6 * tj + ti] = acc; ts.M[6 * ti + tj] = acc; ts.M[6 * tj + ti] = acc; }
      else ts.b[rb] = acc;
    }
    __syncwarp();
    pf_ldlt6_warp(ts, lane);
    double dT[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) dT[i] = ts.dT[i];
    pose = pf_aff_mul(pf_exp_map(dT), pose);   // :1876
    double nm = -1;
#pragma unroll
    for (int i = 0; i < 6; ++i) { const double v = fabs(dT[i]); if (v > nm) nm = v; }
    if (nm <= conv) { reported = it; break; }
    if (it + 1 == max_itr && e_first < e_last) pose = pose0;   // :1886-1892
  }
  pf_inv6_warp(ts, lane);
  return reported;
}

#endif  // __CUDACC__
