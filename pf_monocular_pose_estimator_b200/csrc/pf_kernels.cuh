// pf_kernels.cuh -- sm_100a kernels of the PF hot path.
//
// One persistent kernel (`pf_frames_kernel`) runs whole PF updates: the do/while loop of
// fused propagate -> project -> score passes with a device-side exit test
// (pose_estimator.cpp:535-616), the exact fixed-point weight sum (:627-629), the
// (sum, running-max) scan + stratified resampling in source-centric form (:668-682), the
// most-resampled winner (:685-688) and -- in track mode -- the FP64 tail (Gauss-Newton,
// predictPose).  Two launch shapes share the code:
//   * batch mode  : gridDim = (1, n_trackers): one CTA owns one tracker, block-level scans,
//                   __syncthreads only, any number of frames per launch;
//   * grid mode   : gridDim = (G, 1) cooperative launch: G co-resident CTAs share one filter,
//                   reductions through int64/packed-key atomics + grid.sync(), the scan is a
//                   single-pass decoupled look-back over 1024-particle tiles.
// State is SoA FP32: three float4 planes per pose set (row i of [R|t] in plane i), so every
// particle load/store is a coalesced 16-byte access.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include "pf_tail.cuh"
#include "pf_types.h"

namespace cg = cooperative_groups;

#if defined(__CUDACC__)

// ---------------------------------------------------------------------------------------
// small device helpers
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned pf_ord(float f)
{
  unsigned u = __float_as_uint(f);
  return u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__device__ __forceinline__ float pf_unord(unsigned o)
{
  unsigned u = (o & 0x80000000u) ? (o ^ 0x80000000u) : ~o;
  return __uint_as_float(u);
}

struct pf_red {
  unsigned long long maxkey, minkey_inv, cntkey;
  long long sum;
};

__device__ __forceinline__ pf_red pf_red_identity()
{
  pf_red r; r.maxkey = 0ull; r.minkey_inv = 0ull; r.cntkey = 0ull; r.sum = 0ll; return r;
}

// block-wide reduction of pf_red; result valid in every thread.  red_smem: PF_BLOCK/32 + 1 entries.
__device__ __forceinline__ pf_red pf_block_reduce(pf_red v, pf_red* red_smem)
{
  const unsigned full = 0xFFFFFFFFu;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long a = __shfl_xor_sync(full, v.maxkey, o);
    unsigned long long b = __shfl_xor_sync(full, v.minkey_inv, o);
    unsigned long long c = __shfl_xor_sync(full, v.cntkey, o);
    long long s = __shfl_xor_sync(full, v.sum, o);
    v.maxkey = a > v.maxkey ? a : v.maxkey;
    v.minkey_inv = b > v.minkey_inv ? b : v.minkey_inv;
    v.cntkey = c > v.cntkey ? c : v.cntkey;
    v.sum += s;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();                       // protect red_smem reuse
  if (lane == 0) red_smem[warp] = v;
  __syncthreads();
  pf_red r = red_smem[0];
#pragma unroll
  for (int i = 1; i < PF_BLOCK / 32; ++i) {
    pf_red t = red_smem[i];
    r.maxkey = t.maxkey > r.maxkey ? t.maxkey : r.maxkey;
    r.minkey_inv = t.minkey_inv > r.minkey_inv ? t.minkey_inv : r.minkey_inv;
    r.cntkey = t.cntkey > r.cntkey ? t.cntkey : r.cntkey;
    r.sum += t.sum;
  }
  return r;
}

// Combine the CTA results of one reduction round across the grid (grid mode) and return the
// global result to every thread of every CTA.  Slot rotation: round e uses slot e%3; slot
// (e+2)%3 (last read before the barrier of round e) is re-zeroed by CTA 0 after the barrier.
__device__ __forceinline__ pf_red pf_grid_reduce(pf_red v, pf_red* red_smem, const pf_kparams& p, unsigned& round)
{
  pf_red r = pf_block_reduce(v, red_smem);
  if (p.coop) {
    pf_accum* slot = p.accum + (round % 3u);
    if (threadIdx.x == 0) {
      atomicMax(&slot->maxkey, r.maxkey);
      atomicMax(&slot->minkey_inv, r.minkey_inv);
      atomicMax(&slot->cntkey, r.cntkey);
      atomicAdd((unsigned long long*)&slot->sum, (unsigned long long)r.sum);
    }
    cg::this_grid().sync();
    r.maxkey = *(volatile unsigned long long*)&slot->maxkey;
    r.minkey_inv = *(volatile unsigned long long*)&slot->minkey_inv;
    r.cntkey = *(volatile unsigned long long*)&slot->cntkey;
    r.sum = *(volatile long long*)&slot->sum;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      pf_accum* z = p.accum + ((round + 2u) % 3u);
      z->maxkey = 0ull; z->minkey_inv = 0ull; z->cntkey = 0ull; z->sum = 0ll;
    }
  }
  ++round;
  return r;
}

__device__ __forceinline__ void pf_stamp(const pf_kparams& p, int slot)
{
  if (p.timeline && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.timeline[slot] = t;
  }
}

__device__ __forceinline__ void pf_sync_all(const pf_kparams& p)
{
  if (p.coop) cg::this_grid().sync(); else __syncthreads();
}

__device__ __forceinline__ pf_pose pf_load_pose(const float4* base, int N, int n)
{
  const float4 r0 = base[n], r1 = base[(size_t)N + n], r2 = base[2 * (size_t)N + n];
  pf_pose T;
  T.p[0] = r0.x; T.p[1] = r0.y; T.p[2] = r0.z; T.p[3] = r0.w;
  T.p[4] = r1.x; T.p[5] = r1.y; T.p[6] = r1.z; T.p[7] = r1.w;
  T.p[8] = r2.x; T.p[9] = r2.y; T.p[10] = r2.z; T.p[11] = r2.w;
  return T;
}
__device__ __forceinline__ void pf_store_pose(float4* base, int N, int n, const pf_pose& T)
{
  base[n] = make_float4(T.p[0], T.p[1], T.p[2], T.p[3]);
  base[(size_t)N + n] = make_float4(T.p[4], T.p[5], T.p[6], T.p[7]);
  base[2 * (size_t)N + n] = make_float4(T.p[8], T.p[9], T.p[10], T.p[11]);
}

// Source index of output slot n after a resampled frame: first i with kend[i] > min(n, Ktot-1)
// (two-level search: per-tile ends, then inside the 1024-source tile).  Reproduces :681 incl.
// the "not found keeps the previous index" tail (Q7).
// kend / tile_kend are rewritten by the scan stage of the same launch in multi-frame runs, so
// they are read with plain (coherent) loads, never through the non-coherent path.
__device__ __forceinline__ int pf_source_index(const int* kend, const int* tile_kend, int ntiles, int N, int n)
{
  const int ktot = tile_kend[ntiles - 1];
  int k = n < ktot ? n : ktot - 1;
  if (k < 0) return n;                 // nothing was ever found: the slot keeps its own particle
  int lo = 0, hi = ntiles - 1;         // smallest tile with tile_kend > k
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (tile_kend[mid] > k) hi = mid; else lo = mid + 1;
  }
  int a = lo * PF_SCAN_TILE, b = a + PF_SCAN_TILE - 1;
  if (b > N - 1) b = N - 1;
  while (a < b) {
    int mid = (a + b) >> 1;
    if (kend[mid] > k) b = mid; else a = mid + 1;
  }
  return a;
}

// ---------------------------------------------------------------------------------------
// shared-memory frame context (uniform per CTA)
// ---------------------------------------------------------------------------------------
struct pf_ctx {
  pf_frame_const fc;
  float obj[PF_MAX_LED * 3];
  int src_buf, src_mode, ntiles_scan;
  uint32_t scan_epoch;
};

// ---------------------------------------------------------------------------------------
// Stage 1: fused propagate -> project -> score over this CTA's tiles for iteration `it`
// (:543-605).  first_pass: resolve and cache the resampling source index of every slot.
// ---------------------------------------------------------------------------------------
template <int L>
__device__ __forceinline__ pf_red pf_stage_score(const pf_kparams& p, const pf_ctx& cx, const pf_f2* blobs,
                                                 int trk, int it, bool first_pass)
{
  const int N = p.N, G = gridDim.x;
  const size_t tb = (size_t)trk * 3 * (size_t)N;
  // no __restrict__/const on buffers that other stages of this launch write
  const float4* src = (cx.src_buf == 0 ? p.pose0 : p.pose1) + tb;
  float4* dst = (cx.src_buf == 0 ? p.pose1 : p.pose0) + tb;
  float* w = p.w + (size_t)trk * N;
  int* idx = p.idx + (size_t)trk * N;
  const int* kend = p.kend + (size_t)trk * N;
  const int* tile_kend = p.tile_kend + (size_t)trk * cx.ntiles_scan;

  pf_pose Pm, Cm;
#pragma unroll
  for (int i = 0; i < 12; ++i) { Pm.p[i] = cx.fc.P[i]; Cm.p[i] = cx.fc.C[i]; }
  const bool use_cam = cx.fc.isi > 1;                       // :555-560
  const bool use_pred = use_cam && (it % 10) != 0;
  const float scale = 1.0f + 0.025f * (float)(it / 10);    // :563 (integer division)
  const uint64_t step0 = 12ull * (uint64_t)(N - 2) * (uint64_t)it;

  pf_red acc = pf_red_identity();
  const int ntiles = (N + PF_BLOCK - 1) / PF_BLOCK;
  for (int tile = blockIdx.x; tile < ntiles; tile += G) {
    const int n = tile * PF_BLOCK + threadIdx.x;
    if (n < N) {
      int s = n;
      if (cx.src_mode == 1) {
        if (first_pass) { s = pf_source_index(kend, tile_kend, cx.ntiles_scan, N, n); idx[n] = s; }
        else s = idx[n];
      }
      pf_pose T;
      if (n == 0) {
#pragma unroll
        for (int i = 0; i < 12; ++i) T.p[i] = cx.fc.Tcur[i];      // :547
      } else if (n == 1) {
#pragma unroll
        for (int i = 0; i < 12; ++i) T.p[i] = cx.fc.Tpred[i];     // :551
      } else {
        pf_pose B = pf_load_pose(src, N, s);
        if (use_cam) B = pf_affine_mul(Cm, B);
        if (use_pred) B = pf_affine_mul(B, Pm);
        uint32_t x = pf_rng_skip(cx.fc.seed_state, step0 + 12ull * (uint64_t)(n - 2), p.rng_tab);
        T = pf_propagate(B, x, cx.fc.noise, scale);
      }
      pf_store_pose(dst, N, n, T);
      float u[L], v[L];
#pragma unroll
      for (int j = 0; j < L; ++j) pf_project(p.K9f, T, cx.obj[3 * j], cx.obj[3 * j + 1], cx.obj[3 * j + 2], u[j], v[j]);
      const float wn = pf_score<L, false>(blobs, cx.fc.D, u, v, p.score, nullptr, nullptr);
      w[n] = wn;
      const unsigned o = pf_ord(wn);
      const unsigned long long kmax = ((unsigned long long)o << 32) | (unsigned)(~(unsigned)n);
      const unsigned long long kmin = ~(((unsigned long long)o << 32) | (unsigned)n);
      acc.maxkey = kmax > acc.maxkey ? kmax : acc.maxkey;
      acc.minkey_inv = kmin > acc.minkey_inv ? kmin : acc.minkey_inv;
      acc.sum += pf_w_to_fx(wn);
    }
  }
  return acc;
}

// ---------------------------------------------------------------------------------------
// Stage 3: normalise + scan + stratified resampling in source-centric form (:627-629, :668-682).
// The scan runs on sg*Q.30 weights with the (sum, running max) operator; for every source i
// the cumulative offspring count kend[i] = K(M_i) is stored (4 B) instead of the FP64 prefix,
// count_i = kend[i] - kend[i-1], and the most-resampled particle (:685-686) falls out of the
// same pass.  Batch mode carries the prefix from tile to tile in registers; grid mode uses a
// single-pass decoupled look-back over the tile descriptors.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ pf_sm pf_shfl_sm(const pf_sm& v, int srcLane)
{
  pf_sm r;
  r.s = __shfl_sync(0xFFFFFFFFu, v.s, srcLane);
  r.m = __shfl_sync(0xFFFFFFFFu, v.m, srcLane);
  return r;
}

// executed by warp 0 of the CTA that owns `tile`; returns the exclusive prefix of the tile
__device__ __forceinline__ pf_sm pf_lookback(pf_tile_desc* desc, int tile, uint32_t epoch, const pf_sm& agg)
{
  const int lane = threadIdx.x & 31;
  volatile pf_tile_desc* vd = desc;
  if (lane == 0) {
    vd[tile].agg_s = agg.s;
    vd[tile].agg_m = agg.m;
    __threadfence();
    vd[tile].flag = epoch * 4u + 1u;
  }
  pf_sm excl = pf_sm_identity();
  for (int q0 = tile - 1; q0 >= 0; q0 -= 32) {
    const int q = q0 - lane;
    pf_sm val = pf_sm_identity();
    unsigned st = 2u;                       // tiles before tile 0: inclusive identity
    if (q >= 0) {
      unsigned f;
      do { f = vd[q].flag; } while ((f >> 2) != epoch);
      __threadfence();
      st = f & 3u;
      if (st == 2u) { val.s = vd[q].inc_s; val.m = vd[q].inc_m; }
      else          { val.s = vd[q].agg_s; val.m = vd[q].agg_m; }
    }
    const unsigned ball = __ballot_sync(0xFFFFFFFFu, st == 2u);
    const int first = ball ? (__ffs(ball) - 1) : 31;   // nearest predecessor with an inclusive prefix
    pf_sm wf = pf_sm_identity();
    for (int l = first; l >= 0; --l) wf = pf_sm_combine(wf, pf_shfl_sm(val, l));
    excl = pf_sm_combine(wf, excl);
    if (ball) break;
  }
  if (lane == 0) {
    const pf_sm inc = pf_sm_combine(excl, agg);
    vd[tile].inc_s = inc.s;
    vd[tile].inc_m = inc.m;
    __threadfence();
    vd[tile].flag = epoch * 4u + 2u;
  }
  return excl;
}

__device__ __forceinline__ pf_red pf_stage_scan(const pf_kparams& p, const pf_ctx& cx, int trk, long long S_fx,
                                                int iters_total, pf_sm* sm_warp, pf_sm* sm_bcast)
{
  const int N = p.N, G = gridDim.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* w = p.w + (size_t)trk * N;
  int* kend = p.kend + (size_t)trk * N;
  int* tile_kend = p.tile_kend + (size_t)trk * cx.ntiles_scan;
  const long long sg = S_fx < 0 ? -1ll : 1ll;
  const uint32_t x_base = pf_rng_skip(cx.fc.seed_state, 12ull * (uint64_t)(N - 2) * (uint64_t)iters_total, p.rng_tab);
  const pf_resample_cfg rc = pf_make_resample_cfg(x_base, N, S_fx);

  pf_sm carry = pf_sm_identity();
  pf_red acc = pf_red_identity();
  for (int tile = blockIdx.x; tile < cx.ntiles_scan; tile += G) {
    const int i0 = tile * PF_SCAN_TILE + threadIdx.x * PF_SCAN_ITEMS;
    long long e[PF_SCAN_ITEMS];
    pf_sm a = pf_sm_identity();
#pragma unroll
    for (int j = 0; j < PF_SCAN_ITEMS; ++j) {
      e[j] = (i0 + j < N) ? sg * pf_w_to_fx(w[i0 + j]) : 0ll;
      pf_sm t; t.s = e[j]; t.m = e[j];
      a = pf_sm_combine(a, t);
    }
    // warp inclusive scan of the per-thread aggregates
    pf_sm incl = a;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      pf_sm t;
      t.s = __shfl_up_sync(0xFFFFFFFFu, incl.s, o);
      t.m = __shfl_up_sync(0xFFFFFFFFu, incl.m, o);
      if (lane >= o) incl = pf_sm_combine(t, incl);
    }
    pf_sm lane_excl;
    lane_excl.s = __shfl_up_sync(0xFFFFFFFFu, incl.s, 1);
    lane_excl.m = __shfl_up_sync(0xFFFFFFFFu, incl.m, 1);
    if (lane == 0) lane_excl = pf_sm_identity();
    if (lane == 31) sm_warp[warp] = incl;
    __syncthreads();
    pf_sm warp_excl = pf_sm_identity(), tile_agg = pf_sm_identity();
#pragma unroll
    for (int i = 0; i < PF_BLOCK / 32; ++i) {
      const pf_sm t = sm_warp[i];
      if (i < warp) warp_excl = pf_sm_combine(warp_excl, t);
      tile_agg = pf_sm_combine(tile_agg, t);
    }
    pf_sm tile_excl;
    if (!p.coop) {
      tile_excl = carry;
      carry = pf_sm_combine(carry, tile_agg);
    } else {
      if (warp == 0) {
        const pf_sm ex = pf_lookback(p.desc, tile, cx.scan_epoch, tile_agg);
        if (lane == 0) *sm_bcast = ex;
      }
      __syncthreads();
      tile_excl = *sm_bcast;
    }
    pf_sm pre = pf_sm_combine(pf_sm_combine(tile_excl, warp_excl), lane_excl);
    // cumulative offspring counts of my items
    const long long m_before = pre.m;
    int ke[PF_SCAN_ITEMS];
    long long m_prev = m_before;
    int k_prev = -1;                          // K(m_before), resolved below
#pragma unroll
    for (int j = 0; j < PF_SCAN_ITEMS; ++j) {
      pf_sm t; t.s = e[j]; t.m = e[j];
      pre = pf_sm_combine(pre, t);
      if (j > 0 && pre.m == m_prev) ke[j] = ke[j - 1];
      else ke[j] = pf_count_leq(pre.m, rc, p.rng_tab);
      m_prev = pre.m;
    }
    k_prev = __shfl_up_sync(0xFFFFFFFFu, ke[PF_SCAN_ITEMS - 1], 1);
    if (lane == 0) k_prev = pf_count_leq(m_before, rc, p.rng_tab);
#pragma unroll
    for (int j = 0; j < PF_SCAN_ITEMS; ++j) {
      if (i0 + j < N) {
        kend[i0 + j] = ke[j];
        const unsigned cnt = (unsigned)(ke[j] - k_prev);
        const unsigned long long key = ((unsigned long long)cnt << 32) | (unsigned)(~(unsigned)(i0 + j));
        acc.cntkey = key > acc.cntkey ? key : acc.cntkey;
      }
      k_prev = ke[j];
    }
    if (threadIdx.x == PF_BLOCK - 1) tile_kend[tile] = ke[PF_SCAN_ITEMS - 1];
    __syncthreads();                          // sm_warp / sm_bcast reuse
  }
  return acc;
}

// ---------------------------------------------------------------------------------------
// Tail of one frame, executed by thread 0 of CTA 0 of the tracker: winner, correspondences,
// result record, bookkeeping; in track mode also Gauss-Newton + updatePose + next frame's
// prediction (:685-703, :707-719, :2011-2035, :995-1010).
// ---------------------------------------------------------------------------------------
template <int L>
__device__ __noinline__ void pf_frame_tail(const pf_kparams& p, const pf_ctx& cx, const pf_f2* blobs, int trk, int frame,
                              int iters, int best_it, float highest, const pf_red& kept, bool accept,
                              bool resampled, const pf_red& cnt)
{
  const int N = p.N;
  pf_tracker_state* st = p.state + trk;
  pf_frame_result* res = p.result + (p.mode == 1 ? (size_t)trk * p.total_frames + frame : (size_t)trk);
  const int dst_buf = 1 - cx.src_buf;
  const float4* T = (dst_buf == 0 ? p.pose0 : p.pose1) + (size_t)trk * 3 * (size_t)N;

  const int arg_max = (int)(~(unsigned)(kept.maxkey & 0xFFFFFFFFull));
  const int arg_min = (int)(unsigned)((~kept.minkey_inv) & 0xFFFFFFFFull);
  int q;
  if (resampled) q = (int)(~(unsigned)(cnt.cntkey & 0xFFFFFFFFull));   // :686 first max of counterMeas
  else q = kept.sum < 0 ? arg_min : arg_max;                            // :714 on the normalised weights (Q6)

  res->status = accept ? 1 : 4;
  res->iterations = iters;
  res->best_iteration = best_it;
  res->resampled = resampled ? 1 : 0;
  res->highest_prob = highest;
  res->most_likely_particle_idx = accept ? (best_it >= 0 ? arg_max : -1) : q;
  res->most_likely_pose_idx = resampled ? q : -1;
  res->max_count = resampled ? (unsigned)(cnt.cntkey >> 32) : 0u;
  res->sum_fx = kept.sum;

  const pf_pose Tw = pf_load_pose(T, N, q);
#pragma unroll
  for (int i = 0; i < 12; ++i) res->winner_pose[i] = Tw.p[i];
  unsigned pairs[2 * PF_MAX_LED];
  int npairs = 0;
  {
    float u[L], v[L];
#pragma unroll
    for (int j = 0; j < L; ++j) pf_project(p.K9f, Tw, cx.obj[3 * j], cx.obj[3 * j + 1], cx.obj[3 * j + 2], u[j], v[j]);
    (void)pf_score<L, true>(blobs, cx.fc.D, u, v, p.score, pairs, &npairs);
  }
  for (int i = 0; i < 2 * PF_MAX_LED; ++i) res->pairs[i] = i < 2 * npairs ? pairs[i] : 0u;
  res->npairs = npairs;
  res->gn_iters = -1;
  res->flag_fail = accept ? 1.0f : 4.0f;

  // particle-set bookkeeping for the next frame
  st->t_buf = dst_buf;
  if (resampled) {
    st->src_buf = dst_buf;
    st->src_mode = 1;
    st->scan_epoch = cx.scan_epoch + 1u;
  }
  st->frames_done += 1;
  pf_stamp(p, 7);

  if (p.mode != 1) return;

  // ---- device tail (track mode): optimiseAndUpdatePose + next predictPose -------------
  pf_aff cur, prev;
  for (int i = 0; i < 12; ++i) { cur.a[i] = st->cur[i]; prev.a[i] = st->prev[i]; }
  double t_cur = st->t_cur, t_prev = st->t_prev;
  const double t_pred = st->t_pred;
  uint32_t isi = st->isi;
  if (resampled) {
    pf_aff pose = pf_aff_from_f32(Tw);                 // predicted_pose_ = PoseParticle[q]  (:687)
    const pf_aff before = pose;
    const double* img = p.blobs + ((size_t)trk * p.total_frames + frame) * (size_t)p.D_max * 2;
    double cov[36];
    res->gn_iters = pf_optimise_pose(p.K9d, p.obj_d + (size_t)trk * PF_MAX_LED * 3, img, pairs, npairs, pose, cov);
    if (isi < 2) ++isi;                                // :2027-2030
    prev = cur;                                        // updatePose :2013-2014
    cur = pose;
    if (t_pred - t_cur > 0.001 || t_pred < t_cur) { t_prev = t_cur; t_cur = t_pred; }   // :2016-2020
    float flag = 1.0f;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        if (fabs(before.a[4 * i + j] - pose.a[4 * i + j]) >= 0.3) flag = 1.5f;           // :693-701
    res->flag_fail = flag;
    for (int i = 0; i < 36; ++i) { res->cov[i] = cov[i]; st->cov[i] = cov[i]; }
    for (int i = 0; i < 12; ++i) { st->pred[i] = pose.a[i]; }
  } else {
    isi = 0;                                           // :710
    st->failed = 1;
    for (int i = 0; i < 12; ++i) st->pred[i] = (double)Tw.p[i];   // :716
  }
  for (int i = 0; i < 12; ++i) { st->cur[i] = cur.a[i]; st->prev[i] = prev.a[i]; res->pose[i] = cur.a[i]; }
  st->t_cur = t_cur; st->t_prev = t_prev; st->isi = isi;

  if (resampled && frame + 1 < p.total_frames) {
    const double t_next = p.times[frame + 1];
    pf_aff P, predn;
    pf_predict_pose(prev, cur, t_prev, t_cur, t_next, P, predn);   // :234 -> :995-1010
    st->t_pred = t_next;
    for (int i = 0; i < 12; ++i) st->pred[i] = predn.a[i];
    const size_t fi = (size_t)trk * p.total_frames + frame + 1;
    pf_make_frame_const(p.fconst[trk], cur, predn, P, pf_aff_identity(), isi, t_next - t_cur, p.seeds[fi],
                        p.blob_count[fi], p.min_ang, p.max_ang, p.min_tr, p.max_tr);
  }
}

// ---------------------------------------------------------------------------------------
// The persistent frame kernel.
// ---------------------------------------------------------------------------------------
template <int L>
__global__ void __launch_bounds__(PF_BLOCK, 3) pf_frames_kernel(const __grid_constant__ pf_kparams p)
{
  extern __shared__ __align__(16) unsigned char pf_dyn_smem[];
  pf_f2* blobs = reinterpret_cast<pf_f2*>(pf_dyn_smem);
  __shared__ pf_ctx cx;
  __shared__ pf_red red_smem[PF_BLOCK / 32];
  __shared__ pf_sm sm_warp[PF_BLOCK / 32];
  __shared__ pf_sm sm_bcast;

  const int trk = p.tracker_base + blockIdx.y;
  const int N = p.N;
  pf_tracker_state* st = p.state + trk;
  unsigned round = 0;

  for (int f = 0; f < p.n_frames; ++f) {
    const int frame = p.frame0 + f;
    // ---- frame context (uniform over the tracker's CTAs: written before the last barrier) ----
    if (threadIdx.x == 0) {
      cx.fc = p.fconst[trk];
      cx.src_buf = st->src_buf;
      cx.src_mode = st->src_mode;
      cx.scan_epoch = st->scan_epoch;
      cx.ntiles_scan = (N + PF_SCAN_TILE - 1) / PF_SCAN_TILE;
    }
    if (threadIdx.x < 3 * L) cx.obj[threadIdx.x] = p.obj_f[(size_t)trk * PF_MAX_LED * 3 + threadIdx.x];
    const int failed = (p.mode == 1) ? st->failed : 0;
    pf_stamp(p, 0);
    __syncthreads();
    if (failed) {   // track lost earlier in this sequence: the host must re-initialise (:707-719)
      if (blockIdx.x == 0 && threadIdx.x == 0) p.result[(size_t)trk * p.total_frames + frame].status = 0;
      continue;
    }
    const int D = cx.fc.D;
    {
      const double* bsrc = p.blobs + ((size_t)trk * (p.mode == 1 ? p.total_frames : 1) + (p.mode == 1 ? frame : 0)) *
                                         (size_t)p.D_max * 2;
      for (int i = threadIdx.x; i < D; i += PF_BLOCK) {
        pf_f2 z; z.x = (float)bsrc[2 * i]; z.y = (float)bsrc[2 * i + 1];
        blobs[i] = z;
      }
    }
    __syncthreads();
    pf_stamp(p, 1);

    // ---- do { propagate, project, score } while (iter < 80 && max < L*min(5,numLED))  (:535-616)
    const float thr_exit = (float)(L * (D < 5 ? D : 5));
    float highest = 0.0f;
    int best_it = -1, it = 0;
    pf_red kept = pf_red_identity(), last = pf_red_identity();
    for (;;) {
      pf_red r = pf_stage_score<L>(p, cx, blobs, trk, it, it == 0);
      if (it == 0) pf_stamp(p, 2);
      r = pf_grid_reduce(r, red_smem, p, round);
      if (it == 0) pf_stamp(p, 3);
      const float m = pf_unord((unsigned)(r.maxkey >> 32));
      last = r;
      ++it;
      if (m > highest) { highest = m; best_it = it - 1; kept = r; }   // :608-614 (strict >, from 0)
      if (!(it < PF_MAX_ITER && m < thr_exit)) break;                // :616
    }
    const int iters = it;
    if (best_it < 0) kept = last;                                    // :619: nothing to restore
    else if (best_it != iters - 1) {
      // the kept iteration is not the last one: regenerate its particles and weights (exact,
      // the RNG is addressed by (iteration, particle))
      (void)pf_stage_score<L>(p, cx, blobs, trk, best_it, false);
      pf_sync_all(p);
    }

    // ---- accept test (:633) and resampling (:666-682)
    const bool accept = (kept.sum != 0) && (highest > (float)(L * (D < 3 ? D : 3)));
    const bool resampled = accept && cx.fc.isi > 0;
    pf_red cnt = pf_red_identity();
    if (resampled) {
      pf_stamp(p, 4);
      cnt = pf_stage_scan(p, cx, trk, kept.sum, iters, sm_warp, &sm_bcast);
      pf_stamp(p, 5);
      cnt = pf_grid_reduce(cnt, red_smem, p, round);
      pf_stamp(p, 6);
    }

    if (blockIdx.x == 0 && threadIdx.x == 0)
      pf_frame_tail<L>(p, cx, blobs, trk, frame, iters, best_it, highest, kept, accept, resampled, cnt);
    pf_stamp(p, 8);
    if (f + 1 < p.n_frames) { __threadfence(); pf_sync_all(p); }
    pf_stamp(p, 9);
  }
}

// Materialise newPoseEstimation (the resampled set) for the getters / ROI: out[n] = T[idx(n)].
__global__ void __launch_bounds__(PF_BLOCK) pf_gather_kernel(const float4* T, const int* kend, const int* tile_kend,
                                                             int ntiles, int N, int src_mode, float4* out, int* idx_out)
{
  const int n = blockIdx.x * PF_BLOCK + threadIdx.x;
  if (n >= N) return;
  const int s = src_mode == 1 ? pf_source_index(kend, tile_kend, ntiles, N, n) : n;
  out[n] = T[s];
  out[(size_t)N + n] = T[(size_t)N + s];
  out[2 * (size_t)N + n] = T[2 * (size_t)N + s];
  if (idx_out) idx_out[n] = s;
}

#endif  // __CUDACC__
