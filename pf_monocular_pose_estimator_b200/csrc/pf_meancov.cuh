// pf_meancov.cuh -- K7: weighted mean pose and 6x6 covariance of the particle set.
// An extra asked for by the north star ("reduce to a weighted mean pose and covariance"); the
// reference has no counterpart (its output is the Gauss-Newton pose of the most-resampled particle,
// pose_estimator.cpp:684-690), so this sits beside the frame kernel, behind PFMPE_FLAG_MEAN_COV,
// and is checked against the CPU restatement oracle/pf_oracle.c:orc_weighted_mean_cov.
//
//   w_n  = raw score of the kept iteration / weight sum            (:627-629)
//   tbar = sum w_n t_n,  M = sum w_n R_n,  Rbar = closest rotation to M (U V^T with det +1)
//   xi_n = [t_n - tbar ; log(Rbar^T R_n)^vee],  cov = sum w_n xi_n xi_n^T
//
// Two streaming passes over T (48 B + 4 B per particle each, HBM-bound), FP64 accumulation,
// deterministic: per-block partials are combined in block order by one small block.
#pragma once

#include "pf_tail.cuh"
#include "pf_types.h"

#if defined(__CUDACC__)

#define PF_MC_THREADS 256
#define PF_MC_MAXBLK 592   /* 4 per SM */

// block-wide sum of K doubles per thread; result in out[0..K) (written by the first K threads)
template <int K>
__device__ __forceinline__ void pf_mc_block_sum(double (&v)[K], double* smem /* [8][K] */, double* out)
{
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    double x = v[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) smem[warp * K + j] = x;
  }
  __syncthreads();
  if (threadIdx.x < K) {
    double s = 0;
    for (int wq = 0; wq < PF_MC_THREADS / 32; ++wq) s += smem[wq * K + threadIdx.x];
    out[threadIdx.x] = s;
  }
  __syncthreads();
}

__device__ __forceinline__ const float4* pf_mc_T(const pf_kparams& p, int trk)
{
  const int tb = p.state[trk].t_buf;
  return (tb == 0 ? p.pose0 : p.pose1) + (size_t)trk * 3 * (size_t)p.pstride;
}

// S_all: the weight sum of the WHOLE filter when this handle holds one shard of it (0: this handle's own record)
__device__ __forceinline__ double pf_mc_inv_sum(const pf_kparams& p, int trk, long long S_all)
{
  const long long s = S_all != 0 ? S_all : p.result[trk].sum_fx;
  return s != 0 ? 1073741824.0 / (double)s : 1.0;    // probPart / probPartSum only if the sum is non-zero (:628)
}

// pass 1: partial sums of w t (3) and w R (9, row-major)
__global__ void __launch_bounds__(PF_MC_THREADS) pf_meancov_pass1(const pf_kparams p, double* part, long long S_all)
{
  __shared__ double sm[8 * 12];
  const int trk = p.tracker_base + blockIdx.y, N = p.N;
  const float4* T = pf_mc_T(p, trk);
  const float* w = p.w + (size_t)trk * N;
  const double inv = pf_mc_inv_sum(p, trk, S_all);
  double acc[12];
#pragma unroll
  for (int j = 0; j < 12; ++j) acc[j] = 0;
  for (int n = blockIdx.x * PF_MC_THREADS + threadIdx.x; n < N; n += gridDim.x * PF_MC_THREADS) {
    const float4 r0 = T[n], r1 = T[(size_t)p.pstride + n], r2 = T[2 * (size_t)p.pstride + n];
    const double wn = (double)w[n] * inv;
    acc[0] += wn * r0.x; acc[1] += wn * r0.y; acc[2] += wn * r0.z; acc[3] += wn * r0.w;
    acc[4] += wn * r1.x; acc[5] += wn * r1.y; acc[6] += wn * r1.z; acc[7] += wn * r1.w;
    acc[8] += wn * r2.x; acc[9] += wn * r2.y; acc[10] += wn * r2.z; acc[11] += wn * r2.w;
  }
  pf_mc_block_sum<12>(acc, sm, part + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 12);
}

// closest rotation (det +1) to the 3x3 matrix in rows 0..2 / columns 0..2 of m (row stride 4):
// eigenvectors of M^T M by cyclic Jacobi, u_i = M v_i / sigma_i for the two largest, third columns
// by cross products (det U = det V = +1), R = U V^T.
__host__ __device__ inline void pf_mc_closest_rotation(const double* m, double* R /* row stride 4 */)
{
  double S[3][3], V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) S[i][j] = m[i] * m[j] + m[4 + i] * m[4 + j] + m[8 + i] * m[8 + j];
  for (int sweep = 0; sweep < 40; ++sweep) {
    if (S[0][1] * S[0][1] + S[0][2] * S[0][2] + S[1][2] * S[1][2] < 1e-300) break;
    for (int a = 0; a < 2; ++a)
      for (int b = a + 1; b < 3; ++b) {
        if (S[a][b] == 0) continue;
        const double th = (S[b][b] - S[a][a]) / (2 * S[a][b]);
        const double t = (th >= 0 ? 1.0 : -1.0) / (fabs(th) + sqrt(th * th + 1));
        const double c = 1 / sqrt(t * t + 1), s = t * c;
        for (int k = 0; k < 3; ++k) { const double x = S[k][a], y = S[k][b]; S[k][a] = c * x - s * y; S[k][b] = s * x + c * y; }
        for (int k = 0; k < 3; ++k) { const double x = S[a][k], y = S[b][k]; S[a][k] = c * x - s * y; S[b][k] = s * x + c * y; }
        for (int k = 0; k < 3; ++k) { const double x = V[k][a], y = V[k][b]; V[k][a] = c * x - s * y; V[k][b] = s * x + c * y; }
      }
  }
  int o[3] = {0, 1, 2};                                  // eigenvalues in descending order
  for (int i = 0; i < 2; ++i)
    for (int j = i + 1; j < 3; ++j)
      if (S[o[j]][o[j]] > S[o[i]][o[i]]) { const int t = o[i]; o[i] = o[j]; o[j] = t; }
  double v[3][3], u[3][3];                               // [column][component]
  for (int c = 0; c < 2; ++c)
    for (int k = 0; k < 3; ++k) v[c][k] = V[k][o[c]];
  v[2][0] = v[0][1] * v[1][2] - v[0][2] * v[1][1];
  v[2][1] = v[0][2] * v[1][0] - v[0][0] * v[1][2];
  v[2][2] = v[0][0] * v[1][1] - v[0][1] * v[1][0];
  for (int c = 0; c < 2; ++c) {
    const double ev = S[o[c]][o[c]], sg = sqrt(ev > 0 ? ev : 0);
    for (int r = 0; r < 3; ++r) {
      const double x = m[4 * r] * v[c][0] + m[4 * r + 1] * v[c][1] + m[4 * r + 2] * v[c][2];
      u[c][r] = sg > 0 ? x / sg : 0;
    }
  }
  u[2][0] = u[0][1] * u[1][2] - u[0][2] * u[1][1];
  u[2][1] = u[0][2] * u[1][0] - u[0][0] * u[1][2];
  u[2][2] = u[0][0] * u[1][1] - u[0][1] * u[1][0];
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 3; ++c) R[4 * r + c] = u[0][r] * v[0][c] + u[1][r] * v[1][c] + u[2][r] * v[2][c];
}

// combine the pass-1 partials in block order; mean pose -> result record and scratch
__global__ void pf_meancov_mean(const pf_kparams p, const double* part, int nblk, double* mean /* [tracker][12] */)
{
  __shared__ double m[12];
  const int trk = p.tracker_base + blockIdx.x;
  if (threadIdx.x < 12) {
    double s = 0;
    for (int b = 0; b < nblk; ++b) s += part[((size_t)blockIdx.x * nblk + b) * 12 + threadIdx.x];
    m[threadIdx.x] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double R[12];
    pf_mc_closest_rotation(m, R);
    R[3] = m[3]; R[7] = m[7]; R[11] = m[11];
    for (int i = 0; i < 12; ++i) { mean[(size_t)blockIdx.x * 12 + i] = R[i]; p.result[trk].mean_pose[i] = R[i]; }
  }
}

// pass 2: partial sums of w xi xi^T (21 unique terms, row-major upper triangle)
__global__ void __launch_bounds__(PF_MC_THREADS) pf_meancov_pass2(const pf_kparams p, const double* mean, double* part, long long S_all)
{
  __shared__ double sm[8 * 21];
  __shared__ double mb[12];
  const int trk = p.tracker_base + blockIdx.y, N = p.N;
  if (threadIdx.x < 12) mb[threadIdx.x] = mean[(size_t)blockIdx.y * 12 + threadIdx.x];
  __syncthreads();
  const float4* T = pf_mc_T(p, trk);
  const float* w = p.w + (size_t)trk * N;
  const double inv = pf_mc_inv_sum(p, trk, S_all);
  double acc[21];
#pragma unroll
  for (int j = 0; j < 21; ++j) acc[j] = 0;
  for (int n = blockIdx.x * PF_MC_THREADS + threadIdx.x; n < N; n += gridDim.x * PF_MC_THREADS) {
    const float4 r0 = T[n], r1 = T[(size_t)p.pstride + n], r2 = T[2 * (size_t)p.pstride + n];
    const double Rn[9] = {r0.x, r0.y, r0.z, r1.x, r1.y, r1.z, r2.x, r2.y, r2.z};
    pf_aff D = pf_aff_identity();                        // Rbar^T R_n, zero translation
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) D.a[4 * i + j] = mb[i] * Rn[j] + mb[4 + i] * Rn[3 + j] + mb[8 + i] * Rn[6 + j];
    double lg[6], xi[6];
    pf_log_map(D, lg);
    xi[0] = (double)r0.w - mb[3]; xi[1] = (double)r1.w - mb[7]; xi[2] = (double)r2.w - mb[11];
    xi[3] = lg[3]; xi[4] = lg[4]; xi[5] = lg[5];
    const double wn = (double)w[n] * inv;
    int q = 0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = i; j < 6; ++j) acc[q++] += wn * xi[i] * xi[j];
  }
  pf_mc_block_sum<21>(acc, sm, part + ((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 21);
}

__global__ void pf_meancov_cov(const pf_kparams p, const double* part, int nblk)
{
  __shared__ double c[21];
  const int trk = p.tracker_base + blockIdx.x;
  if (threadIdx.x < 21) {
    double s = 0;
    for (int b = 0; b < nblk; ++b) s += part[((size_t)blockIdx.x * nblk + b) * 21 + threadIdx.x];
    c[threadIdx.x] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int q = 0;
    for (int i = 0; i < 6; ++i)
      for (int j = i; j < 6; ++j) { p.result[trk].mean_cov[6 * i + j] = c[q]; p.result[trk].mean_cov[6 * j + i] = c[q]; ++q; }
  }
}

// Sharded filter ("config 5: allreduce of 33 doubles"): this shard's K partial sums (12 after pass 1, 21 after
// pass 2), combined in block order and sent to every rank's mailbox as the bit patterns of the doubles; the ranks' rows are
// then added in rank order by the frame loop (pfmpe_shard_frame), identically on every rank.
__global__ void pf_shard_mc_publish_kernel(const double* part, int nblk, int K, long long* const* peers, int world, int rank,
                                           int slot, long long seq)
{
  __shared__ long long words[PF_XMAIL_WORDS];
  if ((int)threadIdx.x < K) {
    double s = 0;
    for (int b = 0; b < nblk; ++b) s += part[(size_t)b * K + threadIdx.x];
    words[threadIdx.x] = __double_as_longlong(s);
  }
  __syncthreads();
  if ((int)threadIdx.x < world) {
    long long* dst = peers[threadIdx.x] + ((size_t)slot * world + rank) * PF_XMAIL_WORDS;
    for (int i = 0; i < K; ++i) dst[i] = words[i];
    __threadfence_system();
    *reinterpret_cast<volatile long long*>(dst + PF_XMAIL_WORDS - 1) = seq;
  }
}

#endif  // __CUDACC__
