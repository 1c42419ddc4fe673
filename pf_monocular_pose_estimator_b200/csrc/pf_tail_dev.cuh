// pf_tail_dev.cuh -- warp-cooperative FP64 tail of a tracked frame (device only).
//
// Same semantics as the scalar PF_HD functions of pf_tail.cuh (which the host entry points
// pfmpe_optimise_pose / pfmpe_predict_pose use), restructured for one warp:
//   optimisePose :1805-2009  -> pf_gn_warp: lane q owns correspondence q (projection, residual,
//                               Jacobian rows); 27 lanes own the unique entries of A = sum J^T J and
//                               b = sum J^T e; the 6x6 SPD solve (Eigen's A.ldlt().solve(b), :1872)
//                               is done in 3x3 blocks (Schur complement, cofactor inverses).
//   A.inverse() :2004        -> block inverse from the same factors.
// (file:line are /root/reference/pf_mpe_lib/src/pose_estimator.cpp)
// Everything that is O(1) and scalar (exp/log map, pose products) is evaluated redundantly by
// every lane, so the refined pose is available in registers of all lanes without broadcasts.
#pragma once
#include "pf_tail.cuh"

#if defined(__CUDACC__)

struct pf_tail_smem {
  double J[PF_MAX_LED][14];   // per correspondence: J row 0 (6), J row 1 (6), ex, ey
  double M[36];               // covariance (row-major) after the solve
  double A[36];               // sum J^T J of the last iteration (full symmetric)
  double b[6];
  double pose0[12];           // starting pose (restored if max_itr is hit with a larger error, :1886)
};

#define PF_FULL 0xFFFFFFFFu

__device__ __forceinline__ double pf_shfl_d(double v, int src) { return __shfl_sync(PF_FULL, v, src); }

// ---- O(1) scalar pieces, evaluated redundantly by every lane ---------------------------------
// Same functions as pf_exp_map / pf_log_map / pf_predict_pose (pf_tail.cuh), with the divisions
// of the reference expressions (O/theta*sin, (R-R^T)/f*phi, ...) hoisted into one reciprocal each:
// results agree to a few ulp; the dependent-latency chain is several times shorter.
__device__ __forceinline__ pf_aff pf_exp_map_dev(const double* tw)
{
  const double wx = tw[3], wy = tw[4], wz = tw[5];
  const double th2 = wx * wx + wy * wy + wz * wz;
  double a = 1.0, b = 0.5, c = 1.0 / 6.0;      // R = I + a O + b O^2 ; V = I + b O + c O^2
  if (th2 < 1e-4) {
    // Gauss-Newton steps are tiny: Taylor series in theta^2 (next omitted term < 1e-21 relative);
    // also free of the cancellation in (1 - cos theta) / theta^2 of the closed forms
    a = 1.0 + th2 * (-1.0 / 6.0 + th2 * (1.0 / 120.0 + th2 * (-1.0 / 5040.0 + th2 * (1.0 / 362880.0))));
    b = 0.5 + th2 * (-1.0 / 24.0 + th2 * (1.0 / 720.0 + th2 * (-1.0 / 40320.0 + th2 * (1.0 / 3628800.0))));
    c = 1.0 / 6.0 + th2 * (-1.0 / 120.0 + th2 * (1.0 / 5040.0 + th2 * (-1.0 / 362880.0 + th2 * (1.0 / 39916800.0))));
  } else {
    const double theta = sqrt(th2);
    double st, ct;
    sincos(theta, &st, &ct);
    const double it = 1.0 / theta, it2 = it * it;
    a = st * it; b = (1.0 - ct) * it2; c = (theta - st) * it2 * it;
  }
  const double O[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
  double O2[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) O2[3 * i + j] = O[3 * i] * O[j] + O[3 * i + 1] * O[3 + j] + O[3 * i + 2] * O[6 + j];
  pf_aff T;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    double V[3];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const double id = i == j ? 1.0 : 0.0;
      T.a[4 * i + j] = id + a * O[3 * i + j] + b * O2[3 * i + j];
      V[j] = id + b * O[3 * i + j] + c * O2[3 * i + j];
    }
    T.a[4 * i + 3] = V[0] * tw[0] + V[1] * tw[1] + V[2] * tw[2];
  }
  return T;
}

__device__ __forceinline__ void pf_log_map_dev(const pf_aff& T, double* xi)
{
  const double* a = T.a;
  double dev = 0, nrm = 0;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const double e = a[4 * i + j] - (i == j ? 1.0 : 0.0);
      dev += e * e;
      nrm += a[4 * i + j] * a[4 * i + j];
    }
  double k = 0.0;                                  // w_hat = (R - R^T) * k,  k = phi / (2 sin phi)
  if (!(dev <= 1e-20 * (nrm < 3.0 ? nrm : 3.0))) {
    double c = (a[0] + a[5] + a[10] - 1) / 2;
    c = c > 1 ? 1 : (c < -1 ? -1 : c);
    const double phi = acos(c);
    if (phi != 0) k = phi / (2 * sin(phi));
  }
  const double w0 = (a[9] - a[6]) * k, w1 = (a[2] - a[8]) * k, w2 = (a[4] - a[1]) * k;
  const double wh[9] = {0, -w2, w1, w2, 0, -w0, -w1, w0, 0};
  const double wn = sqrt(w0 * w0 + w1 * w1 + w2 * w2);
  const double t[3] = {a[3], a[7], a[11]};
  const bool t_zero = (t[0] * t[0] + t[1] * t[1] + t[2] * t[2]) <= 0.0;
  double f = 0.0, h = 0.0, d = 0.0;                // A_inv = d I - h w_hat + f w_hat^2
  if (!t_zero) {
    d = 1.0;
    double sw, cw;
    sincos(wn, &sw, &cw);
    if (!(wn == 0 || sw == 0)) { h = 0.5; f = (2 * sw - wn * (1 + cw)) / (2 * wn * wn * sw); }
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    double r = 0.0;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const double w2ij = wh[3 * i] * wh[j] + wh[3 * i + 1] * wh[3 + j] + wh[3 * i + 2] * wh[6 + j];
      const double Aij = (i == j ? d : 0.0) - h * wh[3 * i + j] + f * w2ij;
      r += Aij * t[j];
    }
    xi[i] = r;
  }
  xi[3] = w0; xi[4] = w1; xi[5] = w2;
}

__device__ __forceinline__ void pf_predict_pose_dev(const pf_aff& prev, const pf_aff& cur, double t_prev, double t_cur,
                                                    double t_pred, pf_aff& P, pf_aff& pred)
{
  double d[6];
  pf_log_map_dev(pf_aff_mul(pf_aff_inverse(prev), cur), d);
#pragma unroll
  for (int i = 0; i < 6; ++i) d[i] = d[i] / (t_cur - t_prev) * (t_pred - t_cur);
  P = pf_exp_map_dev(d);
  pred = pf_aff_mul(cur, P);
}

// (the 6x6 SPD solve in 3x3 blocks -- pf_blk6_factor / pf_blk6_solve / pf_blk6_inverse -- lives in pf_tail.cuh: the host form
// of the Gauss-Newton uses it too)

// Correspondences of one pose by the whole warp (calculateEstimationProbability :2385-2445 for the
// winner): lane i scans blobs i, i+32, ...; per LED the (distance, blob index) minimum is combined
// across lanes on a packed key (squared distances are >= 0, so their bit patterns order like the
// values; ties go to the smaller blob index as in the per-particle loop).
template <int L, typename Blob>
__device__ __forceinline__ float pf_score_warp(const Blob* blobs, int D, const float* K9f, bool kstd, const float* dist, const pf_pose& T,
                                               const float* obj, const pf_score_cfg& sc, unsigned* pairs, int* npairs,
                                               int lane)
{
  // rolled over the LEDs (once-per-frame code: small beats fast); lane j ends up holding LED j's key
  unsigned long long mine = ((unsigned long long)__float_as_uint(INFINITY) << 32);
#pragma unroll 1
  for (int j = 0; j < L; ++j) {
    float u, v;
    pf_project(K9f, kstd, T, obj[3 * j], obj[3 * j + 1], obj[3 * j + 2], u, v);
    if (dist) pf_distort(K9f, dist, u, v);
    unsigned long long key = ((unsigned long long)__float_as_uint(INFINITY) << 32);
    for (int i = lane; i < D; i += 32) {
      const Blob z = blobs[i];                            // any order: the key carries the blob index
      const float dx = z.x - u, dy = z.y - v;
      const float d2 = fmaf(dy, dy, dx * dx);
      const unsigned long long k = ((unsigned long long)__float_as_uint(d2) << 32) | (unsigned)z.idx;
      key = k < key ? k : key;
    }
#pragma unroll 1
    for (int o = 16; o > 0; o >>= 1) {
      const unsigned long long t = __shfl_xor_sync(PF_FULL, key, o);
      key = t < key ? t : key;
    }
    if (lane == j) mine = key;
  }
  float m[L];
  int am[L];
#pragma unroll
  for (int j = 0; j < L; ++j) {
    const unsigned long long key = __shfl_sync(PF_FULL, mine, j);
    m[j] = __uint_as_float((unsigned)(key >> 32));
    am[j] = (int)(unsigned)(key & 0xFFFFFFFFull);
  }
  return pf_greedy<L, true>(m, am, D, sc, pairs, npairs);
}

// lower-triangle lane map: lane l < 21 -> (i, j), j <= i
__device__ __forceinline__ void pf_tri_ij(int l, int& i, int& j)
{
  i = 0;
#pragma unroll
  for (int r = 1; r < 6; ++r) if (l >= r * (r + 1) / 2) i = r;
  j = l - i * (i + 1) / 2;
}

// optimisePose (:1805-2009) by one warp.  K9: row-major K (kernel parameter space); obj: L x 3
// marker coordinates (shared memory); img: detections of the frame (D x 2 doubles, global);
// pairs: (LED+1, blob+1) rows in registers.  pose in/out (replicated in every lane).  The
// covariance (:2004) is left in ts.M (row-major).  Returns PubData.numIter (-1 when max_itr is hit).
// cap > 0: give up after `cap` iterations and return -2 (the pose is then an intermediate iterate): the caller hands the
// object to the host, which runs the whole optimisation again with the reference's full 500-iteration semantics -- an
// optimisation that creeps just above 1e-13 for all 500 iterations costs one warp 0.65 ms but a host core 0.2 ms.
template <int L>
__device__ __forceinline__ int pf_gn_warp(const double* K9, const double* obj, const double* img, const unsigned* pairs,
                                          int npairs, pf_aff& pose, pf_tail_smem& ts, int lane, int cap = 0)
{
  const double conv = 1e-13;
  const int max_itr = 500;
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < 12; ++i) ts.pose0[i] = pose.a[i];
  }
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
    if (cap > 0 && it == cap) { reported = -2; break; }
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
      const double ip2 = 1.0 / p2, iz = 1.0 / z;
      const double ex = zx - p0 * ip2, ey = zy - p1 * ip2;
      if (it == 0 || it + 1 == max_itr) en = sqrt(ex * ex + ey * ey);
      const double xz = x * iz, yz = y * iz;                    // computeJacobian (:2163-2192)
      Jr[0] = fx * iz;  Jr[2] = -fx * xz * iz;  Jr[3] = -fx * xz * yz;  Jr[4] = fx * (1.0 + xz * xz);  Jr[5] = -fx * yz;
      Jr[7] = fy * iz;  Jr[8] = -fy * yz * iz;  Jr[9] = -fy * (1.0 + yz * yz);  Jr[10] = fy * xz * yz;  Jr[11] = fy * xz;
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
      if (own_A) { ts.A[6 * ti + tj] = acc; ts.A[6 * tj + ti] = acc; }
      else ts.b[rb] = acc;
    }
    __syncwarp();
    double dT[6];
    {
      pf_blk6 f;
      pf_blk6_factor(ts.A, f);
      pf_blk6_solve(f, ts.A, ts.b, dT);
    }
    pose = pf_aff_mul(pf_exp_map_dev(dT), pose);   // :1876
    double nm = -1;
#pragma unroll
    for (int i = 0; i < 6; ++i) { const double v = fabs(dT[i]); if (v > nm) nm = v; }
    if (nm <= conv) { reported = it; break; }
    if (it + 1 == max_itr && e_first < e_last) {   // :1886-1892
#pragma unroll
      for (int i = 0; i < 12; ++i) pose.a[i] = ts.pose0[i];
    }
  }
  {
    pf_blk6 f;
    pf_blk6_factor(ts.A, f);
    double cov[36];
    pf_blk6_inverse(f, cov);
    __syncwarp();
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 36; ++i) ts.M[i] = cov[i];
    }
    __syncwarp();
  }
  return reported;
}

#endif  // __CUDACC__
