// pf_roi.cuh -- K8: ROI prediction for the LED detector.
// predictMarkerPositionsInImage (pose_estimator.cpp:1036-1052) projects every LED through every
// resampled particle, camMoveInv * newPoseEstimation[j] * predictionMatrix, and through
// predicted_pose_; LEDDetector::determineROI (led_detector.cpp:221-244) then takes the extreme
// values (x_min, y_min from +INF; x_max, y_max from 0; strict comparisons, so a NaN never wins).
// That is the other O(N*L) step of the tracker per frame; here it is one streaming pass over the
// particle set where it lives (through idx[], i.e. without materialising newPoseEstimation) and a
// 16-byte result instead of a device->host copy of all particles.  The two corner points it
// returns go through determineROI's distortion / border / clamping on the host unchanged.
#pragma once

#include "pf_kernels.cuh"

#if defined(__CUDACC__)

__global__ void pf_roi_init_kernel(unsigned* keys, int n_trackers)
{
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_trackers) return;
  keys[4 * t + 0] = pf_ord(INFINITY); keys[4 * t + 1] = pf_ord(INFINITY);
  keys[4 * t + 2] = pf_ord(0.0f);     keys[4 * t + 3] = pf_ord(0.0f);
}

template <int L>
__global__ void __launch_bounds__(PF_BLOCK) pf_roi_kernel(const pf_kparams p, const pf_pose Cm, const pf_pose Pm,
                                                          const pf_pose pred, unsigned* keys)
{
  __shared__ float red[4][PF_BLOCK / 32];
  const int trk = p.tracker_base + blockIdx.y, N = p.N;
  const pf_tracker_state* st = p.state + trk;
  const float4* src = (st->src_buf == 0 ? p.pose0 : p.pose1) + (size_t)trk * 3 * (size_t)p.pstride;
  const int* idx = p.lead ? p.lead : p.idx + (size_t)trk * N;   // sharded filter: lead[] (pf_shard_scatter_kernel)
  const bool through_idx = st->src_mode == 1;
  const bool kstd = p.kstd != 0;
  float obj[3 * L];
#pragma unroll
  for (int i = 0; i < 3 * L; ++i) obj[i] = p.obj_f[(size_t)trk * PF_MAX_LED * 3 + i];
  float x0 = INFINITY, y0 = INFINITY, x1 = 0.0f, y1 = 0.0f;
  for (int n = blockIdx.x * PF_BLOCK + threadIdx.x; n <= N; n += gridDim.x * PF_BLOCK) {
    pf_pose T;
    if (n < N) {
      const int s = through_idx ? idx[n] : n;
      T = pf_affine_mul(pf_affine_mul(Cm, pf_load_pose(src, p.pstride, s)), Pm);
    } else T = pred;                                  // :1047-1050
#pragma unroll
    for (int j = 0; j < L; ++j) {
      float u, v;
      pf_project(p.K9f, kstd, T, obj[3 * j], obj[3 * j + 1], obj[3 * j + 2], u, v);
      if (u < x0) x0 = u;
      if (u > x1) x1 = u;
      if (v < y0) y0 = v;
      if (v > y1) y1 = v;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    x0 = fminf(x0, __shfl_xor_sync(PF_FULL, x0, o)); y0 = fminf(y0, __shfl_xor_sync(PF_FULL, y0, o));
    x1 = fmaxf(x1, __shfl_xor_sync(PF_FULL, x1, o)); y1 = fmaxf(y1, __shfl_xor_sync(PF_FULL, y1, o));
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { red[0][warp] = x0; red[1][warp] = y0; red[2][warp] = x1; red[3][warp] = y1; }
  __syncthreads();
  if (threadIdx.x < 4) {
    float r = red[threadIdx.x][0];
    for (int wq = 1; wq < PF_BLOCK / 32; ++wq) r = threadIdx.x < 2 ? fminf(r, red[threadIdx.x][wq]) : fmaxf(r, red[threadIdx.x][wq]);
    unsigned* k = keys + 4 * (size_t)trk + threadIdx.x;
    if (threadIdx.x < 2) atomicMin(k, pf_ord(r)); else atomicMax(k, pf_ord(r));
  }
}

__global__ void pf_roi_finish_kernel(const unsigned* keys, float* out, int n)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = pf_unord(keys[i]);
}

#endif  // __CUDACC__
