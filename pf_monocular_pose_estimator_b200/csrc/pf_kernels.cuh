// pf_kernels.cuh -- sm_100a kernels of the PF hot path.
//
// One persistent kernel (`pf_frames_kernel`) runs whole PF updates.  Per frame:
//   stage A  do { fused propagate -> project -> score pass } while (...)   (pose_estimator.cpp:535-616)
//            with the device-side exit test; every pass also leaves the exact fixed-point weight
//            sum (:627-629) and, per warp, the (sum, running max, running min) aggregate of the
//            prefix of its particles -- the carry-in of the resampling stage;
//   ---- grid-wide reduction/barrier 1 (max, first argmax, argmin, sum) ----
//   stage B  normalise + prefix + stratified resampling in source-centric form (:668-682) and the
//            most-resampled particle (:685-686);
//   ---- grid-wide reduction/barrier 2 (max offspring count, first index) ----
//   stage C  tail, executed redundantly by warp 0 of every CTA (no third barrier): winner pose,
//            correspondences, result record; in track mode also Gauss-Newton, updatePose, the jump
//            check and the next frame's predictPose (:685-703, :1805-2009, :995-1010), all FP64.
//
// Work decomposition: the particle range is cut into one contiguous chunk per CTA and one
// contiguous sub-chunk per warp, walked 32 particles at a time.  Because a warp's particles are
// contiguous, prefix state is a register carry: there is no __syncthreads inside a stage, no tile
// descriptors and no look-back protocol.  The only cross-CTA prefix step -- the exclusive
// combination of the per-CTA aggregates -- happens after barrier 1, when all of them are final.
// Two launch shapes share the code:
//   * batch mode : gridDim = (1, n_trackers): one CTA owns one tracker (barriers = __syncthreads);
//   * grid mode  : gridDim = (G, 1) cooperative launch: G co-resident CTAs share one filter,
//                  reductions through int64 / packed-key atomics + grid.sync().
// State is SoA FP32: three float4 planes per pose set (row i of [R|t] in plane i), so every
// particle load/store is a coalesced 16-byte access.
#pragma once
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include "pf_tail_dev.cuh"
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

// Once-per-pass glue below is latency-critical single-warp code: loops are kept rolled (#pragma
// unroll 1) on purpose -- its cost is instruction fetch (L0 I-cache ~6 KB, L1.5 32 KB), not issue.

// ordered combination of the warp aggregates of this CTA, by one warp; result in lane 0
__device__ __forceinline__ pf_sm3 pf_warp0_combine_agg(const pf_sm3* warp_agg, int lane)
{
  const int nwarps = blockDim.x >> 5;
  pf_sm3 a = lane < nwarps ? warp_agg[lane] : pf_sm3_identity();
#pragma unroll 1
  for (int o = 1; o < 32; o <<= 1) {
    pf_sm3 t;
    t.s = __shfl_down_sync(PF_FULL, a.s, o); t.mx = __shfl_down_sync(PF_FULL, a.mx, o); t.mn = __shfl_down_sync(PF_FULL, a.mn, o);
    if (lane + o < 32) a = pf_sm3_combine(a, t);
  }
  return a;
}

// Grid-wide barrier (grid mode; all CTAs are co-resident: cooperative launch).  One monotonically
// increasing 64-bit arrival counter that is never reset: barrier number b (counted across
// launches, kept in pf_bar) completes when the counter reaches (b + 1) * gridDim.x.  Arrival is a
// release reduction (no return value to wait for), the wait is an acquire poll by one thread.
struct pf_bar { unsigned long long done; unsigned round; };

// arrive + wait, called by ONE thread after a __syncthreads that follows every thread's writes
__device__ __forceinline__ void pf_bar_arrive_wait(const pf_kparams& p, const pf_bar& bar)
{
  unsigned long long* c = &p.accum[3].cntkey;
  const unsigned long long target = (bar.done + 1ull) * gridDim.x;
  asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(c), "l"(1ull) : "memory");
  unsigned long long v;
  do {
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(c) : "memory");
  } while (v < target);
}

__device__ __forceinline__ void pf_grid_barrier(const pf_kparams& p, pf_bar& bar)
{
  __syncthreads();
  if (threadIdx.x == 0) {
    pf_bar_arrive_wait(p, bar);
    __threadfence();
  }
  __syncthreads();
  bar.done += 1ull;
}

// Reduce pf_red over the CTA and (grid mode) over the grid; result valid in every thread.  Two
// __syncthreads in total: warps -> shared memory -> warp 0 (shuffle reduction; lane 0 publishes the
// CTA's prefix aggregate, adds the CTA result to the round's slot, passes the grid barrier and reads
// the slot back: one reader per CTA) -> shared memory -> everyone.
// Slot rotation: round e uses slot e%3; slot (e+2)%3 (last read before the barrier of round e) is
// re-zeroed by CTA 0 after the barrier.  red_smem: PF_MAX_WARPS + 1 entries.
// Cluster-pair mode (p.coop == 2: the two CTAs of a thread-block cluster share one filter): the CTA results and prefix
// aggregates are exchanged through distributed shared memory -- each CTA stores its entry into both CTAs' xch[] -- and
// one cluster barrier replaces the global atomics and the polling barrier.  xch is double-buffered by round parity: a
// CTA that runs ahead writes the next round's buffer while its partner may still be reading this round's.
struct pf_xch { pf_red r; pf_sm3 agg; };

__device__ __noinline__ pf_red pf_grid_reduce(pf_red v, pf_red* red_smem, const pf_kparams& p, pf_bar& bar,
                                              const pf_sm3* warp_agg, pf_xch (*xch)[2])
{
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll 1
  for (int o = 16; o > 0; o >>= 1) {
    unsigned long long a = __shfl_xor_sync(PF_FULL, v.maxkey, o);
    unsigned long long b = __shfl_xor_sync(PF_FULL, v.minkey_inv, o);
    unsigned long long c = __shfl_xor_sync(PF_FULL, v.cntkey, o);
    long long s = __shfl_xor_sync(PF_FULL, v.sum, o);
    v.maxkey = a > v.maxkey ? a : v.maxkey;
    v.minkey_inv = b > v.minkey_inv ? b : v.minkey_inv;
    v.cntkey = c > v.cntkey ? c : v.cntkey;
    v.sum += s;
  }
  if (lane == 0) red_smem[warp] = v;
  __syncthreads();
  if (warp == 0) {
    const int nwarps = blockDim.x >> 5;
    pf_red r = lane < nwarps ? red_smem[lane] : pf_red_identity();
#pragma unroll 1
    for (int o = 16; o > 0; o >>= 1) {
      unsigned long long a = __shfl_xor_sync(PF_FULL, r.maxkey, o);
      unsigned long long b = __shfl_xor_sync(PF_FULL, r.minkey_inv, o);
      unsigned long long c = __shfl_xor_sync(PF_FULL, r.cntkey, o);
      long long s = __shfl_xor_sync(PF_FULL, r.sum, o);
      r.maxkey = a > r.maxkey ? a : r.maxkey;
      r.minkey_inv = b > r.minkey_inv ? b : r.minkey_inv;
      r.cntkey = c > r.cntkey ? c : r.cntkey;
      r.sum += s;
    }
    if (p.coop == 2) {
      pf_sm3 agg = pf_sm3_identity();
      if (warp_agg) agg = pf_warp0_combine_agg(warp_agg, lane);          // valid in lane 0
      agg.s = __shfl_sync(PF_FULL, agg.s, 0); agg.mx = __shfl_sync(PF_FULL, agg.mx, 0); agg.mn = __shfl_sync(PF_FULL, agg.mn, 0);
      if (lane < 2) {                                     // lane d delivers this CTA's entry to CTA d of the pair (r: all lanes hold it)
        cg::cluster_group cl = cg::this_cluster();
        pf_xch* dst = cl.map_shared_rank(&xch[bar.round & 1u][0], lane);
        pf_xch e;
        e.r = r; e.agg = agg;
        dst[cl.block_rank()] = e;
      }
    } else if (p.coop == 1) {
      pf_sm3 agg = pf_sm3_identity();
      if (warp_agg) agg = pf_warp0_combine_agg(warp_agg, lane);
      if (lane == 0) {
        if (warp_agg) { pf_cta_agg* g = p.cta_agg + blockIdx.x; g->s = agg.s; g->mx = agg.mx; g->mn = agg.mn; }
        pf_accum* slot = p.accum + (bar.round % 3u);
        atomicMax(&slot->maxkey, r.maxkey);
        atomicMax(&slot->minkey_inv, r.minkey_inv);
        atomicMax(&slot->cntkey, r.cntkey);
        atomicAdd((unsigned long long*)&slot->sum, (unsigned long long)r.sum);
        pf_bar_arrive_wait(p, bar);
        r.maxkey = *(volatile const unsigned long long*)&slot->maxkey;
        r.minkey_inv = *(volatile const unsigned long long*)&slot->minkey_inv;
        r.cntkey = *(volatile const unsigned long long*)&slot->cntkey;
        r.sum = *(volatile const long long*)&slot->sum;
        if (blockIdx.x == 0) {
          pf_accum* z = p.accum + ((bar.round + 2u) % 3u);
          z->maxkey = 0ull; z->minkey_inv = 0ull; z->cntkey = 0ull; z->sum = 0ll;
        }
        __threadfence();
      }
    }
    if (lane == 0) red_smem[PF_MAX_WARPS] = r;
  }
  if (p.coop == 2) {
    cg::this_cluster().sync();                            // both entries of this round are in place (release / acquire)
    const pf_xch* x = xch[bar.round & 1u];
    pf_red r;
    r.maxkey = x[0].r.maxkey > x[1].r.maxkey ? x[0].r.maxkey : x[1].r.maxkey;
    r.minkey_inv = x[0].r.minkey_inv > x[1].r.minkey_inv ? x[0].r.minkey_inv : x[1].r.minkey_inv;
    r.cntkey = x[0].r.cntkey > x[1].r.cntkey ? x[0].r.cntkey : x[1].r.cntkey;
    r.sum = x[0].r.sum + x[1].r.sum;
    bar.round += 1u;
    return r;
  }
  __syncthreads();
  if (p.coop == 1) bar.done += 1ull;
  bar.round += 1u;
  return red_smem[PF_MAX_WARPS];
}

__device__ __forceinline__ void pf_stamp(const pf_kparams& p, int slot)
{
  if (p.timeline && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.timeline[slot] = t;
  }
}

__device__ __forceinline__ void pf_sync_all(const pf_kparams& p, pf_bar& bar)
{
  if (p.coop == 1) pf_grid_barrier(p, bar); else if (p.coop == 2) cg::this_cluster().sync(); else __syncthreads();
}

// base: three planes of float4 (row i of [R|t] in plane i), S = plane stride in float4 (= particles per tracker; twice that
// for a sharded filter, whose planes carry a receive area behind the shard's own rows)
__device__ __forceinline__ void pf_prefetch_pose(const float4* base, int S, int n)
{
  asm volatile("prefetch.global.L1 [%0];" ::"l"(base + n));
  asm volatile("prefetch.global.L1 [%0];" ::"l"(base + (size_t)S + n));
  asm volatile("prefetch.global.L1 [%0];" ::"l"(base + 2 * (size_t)S + n));
}

__device__ __forceinline__ pf_pose pf_load_pose(const float4* base, int S, int n)
{
  const float4 r0 = base[n], r1 = base[(size_t)S + n], r2 = base[2 * (size_t)S + n];
  pf_pose T;
  T.p[0] = r0.x; T.p[1] = r0.y; T.p[2] = r0.z; T.p[3] = r0.w;
  T.p[4] = r1.x; T.p[5] = r1.y; T.p[6] = r1.z; T.p[7] = r1.w;
  T.p[8] = r2.x; T.p[9] = r2.y; T.p[10] = r2.z; T.p[11] = r2.w;
  return T;
}
__device__ __forceinline__ void pf_store_pose(float4* base, int S, int n, const pf_pose& T)
{
  base[n] = make_float4(T.p[0], T.p[1], T.p[2], T.p[3]);
  base[(size_t)S + n] = make_float4(T.p[4], T.p[5], T.p[6], T.p[7]);
  base[2 * (size_t)S + n] = make_float4(T.p[8], T.p[9], T.p[10], T.p[11]);
}

// ---------------------------------------------------------------------------------------
// warp-level primitives on fixed-point weights
// ---------------------------------------------------------------------------------------
// sum over the warp of non-negative values < 2^40 (two integer warp reductions)
__device__ __forceinline__ long long pf_warp_sum_nonneg(long long e)
{
  const unsigned lo = (unsigned)(e & 0xFFFFFll), hi = (unsigned)(e >> 20);
  const unsigned slo = __reduce_add_sync(PF_FULL, lo), shi = __reduce_add_sync(PF_FULL, hi);
  return (long long)slo + ((long long)shi << 20);
}

__device__ __forceinline__ long long pf_shfl_ll(long long v, int src) { return __shfl_sync(PF_FULL, v, src); }

// ---------------------------------------------------------------------------------------
// Detections of the frame in shared memory: candidate lists per image column.
//
// The score (:2385-2445) depends, per LED, only on the nearest blob if it lies within the gate
// tol_PF (the greedy loop stops at the first LED whose nearest blob is farther, :2414/:2438).  A
// blob within the gate has |dx| <= tol_PF, so the search is restricted to blobs whose x lies within
// r of the projected LED, r slightly above tol_PF.  The image is cut into at most PF_NCOL columns of
// width >= 1.01 r over [min x - r, max x + r]; column c lists (cand[desc.start .. desc.end)) every
// blob whose interval [x - r, x + r] touches it -- at most three columns per blob -- in ascending
// blob index, as full (x, y, index) records, so that a query is: one column number (an FFMA, a
// saturating conversion, a min), one descriptor load, and a strict-< scan of typically one or two
// records.  pf_col() is a monotone function of x (one correctly rounded FMA, truncation, clamp),
// which is all the exactness argument needs: |u - x_i| <= tol_PF < r  =>  x_i - r <= u <= x_i + r
// =>  col(x_i - r) <= col(u) <= col(x_i + r)  =>  blob i is listed in col(u).  Same minimum, same
// index (ties go to the smaller blob index: lists are ascending, the comparison is strict) as the
// reference's full scan whenever the result matters.  NaN coordinates fall into column 0 and never
// win a comparison.
// ---------------------------------------------------------------------------------------
#define PF_NCOL 128
#define PF_CAND_PER_BLOB 3
struct pf_blob { float x, y; int idx; int pad; };          // 16 B: one LDS.128
struct pf_blobtab {
  float kx, c0, r;
  int D, ncol;
  unsigned desc[PF_NCOL];                                  // (count << 16) | start into cand[]; during the build: counts
  unsigned start[PF_NCOL];                                 // build scratch
};
// cand[0] is a sentinel record at infinity: an empty column points to it with count 0, so that a query can load and
// test "the first candidate" unconditionally (distance +inf) and only loops for a second, third, ... candidate.
// dynamic shared memory of the frame kernel: blobs[D_max] (index order) | cand[1 + 3 D_max] | span[D_max] (u16)
__host__ __device__ __forceinline__ size_t pf_blob_smem_bytes(int d_max)
{
  return (size_t)d_max * (sizeof(pf_blob) * (1 + PF_CAND_PER_BLOB)) + sizeof(pf_blob) + (((size_t)d_max * 2 + 15) & ~(size_t)15);
}

__device__ __forceinline__ unsigned pf_col(const pf_blobtab& bt, float x)
{
  const unsigned c = __float2uint_rz(fmaf(x, bt.kx, bt.c0));   // saturating: negative / NaN -> 0, huge -> 2^32 - 1
  return min(c, (unsigned)(bt.ncol - 1));
}

// executed by ALL threads of the CTA (contains __syncthreads); bsrc: D x 2 doubles in global memory.
// Counting sort by column, parallel over the blobs: every blob adds itself to the counters of the (at most three)
// columns it touches, warp 0 turns the counters into list starts, and every (blob, column) pair finds its place in the
// list -- ascending blob index -- as the number of lower-numbered blobs that touch the same column.
__device__ __noinline__ void pf_stage_blobs(const double* bsrc, int D, float tol_pf, pf_blobtab& bt, pf_blob* blobs, int d_max)
{
  pf_blob* cand = blobs + d_max;
  unsigned short* span = reinterpret_cast<unsigned short*>(cand + 1 + (size_t)PF_CAND_PER_BLOB * d_max);
  const int tid = threadIdx.x, nt = blockDim.x;
  if (tid == 0) { pf_blob z; z.x = INFINITY; z.y = INFINITY; z.idx = 0; z.pad = 0; cand[0] = z; }
  for (int c = tid; c < PF_NCOL; c += nt) bt.desc[c] = 0u;
  for (int i = tid; i < D; i += nt) {
    pf_blob b;
    b.x = (float)bsrc[2 * i]; b.y = (float)bsrc[2 * i + 1]; b.idx = i; b.pad = 0;
    blobs[i] = b;
  }
  __syncthreads();
  if (tid < 32) {
    float lo = INFINITY, hi = -INFINITY;
    for (int i = tid; i < D; i += 32) { const float x = blobs[i].x; lo = fminf(lo, x); hi = fmaxf(hi, x); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      lo = fminf(lo, __shfl_xor_sync(PF_FULL, lo, o));
      hi = fmaxf(hi, __shfl_xor_sync(PF_FULL, hi, o));
    }
    if (tid == 0) {
      const float r = tol_pf * 1.0001f + 1.0e-3f;
      const float span_x = (hi - lo) + 2.0f * r;
      int ncol = 1;
      float kx = 0.0f, c0 = 0.0f;
      if (D > 0 && span_x > 0.0f && span_x < 3.0e38f && r > 0.0f) {
        const float q = span_x / (1.01f * r);
        ncol = q < (float)PF_NCOL ? (int)q : PF_NCOL;
        if (ncol < 1) ncol = 1;
        kx = (float)ncol / span_x;
        c0 = -(lo - r) * kx;
      }
      bt.kx = kx; bt.c0 = c0; bt.r = r; bt.D = D; bt.ncol = ncol;
    }
  }
  __syncthreads();
  for (int i = tid; i < D; i += nt) {
    const float x = blobs[i].x;
    const unsigned clo = pf_col(bt, x - bt.r), chi = pf_col(bt, x + bt.r);
    span[i] = (unsigned short)(clo | (chi << 8));
    for (unsigned c = clo; c <= chi && c < clo + PF_CAND_PER_BLOB; ++c) atomicAdd(&bt.desc[c], 1u);
  }
  __syncthreads();
  if (tid < 32) {
    int h[PF_NCOL / 32], sum = 0;
#pragma unroll
    for (int q = 0; q < PF_NCOL / 32; ++q) { h[q] = (int)bt.desc[tid * (PF_NCOL / 32) + q]; sum += h[q]; }
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(PF_FULL, inc, o);
      if (tid >= o) inc += t;
    }
    int run = inc - sum + 1;                                 // lists start behind the sentinel
#pragma unroll
    for (int q = 0; q < PF_NCOL / 32; ++q) {
      const int c = tid * (PF_NCOL / 32) + q;
      bt.start[c] = (unsigned)run;
      bt.desc[c] = h[q] > 0 ? (((unsigned)h[q] << 16) | (unsigned)run) : 0u;
      run += h[q];
    }
  }
  __syncthreads();
  for (int t = tid; t < PF_CAND_PER_BLOB * D; t += nt) {
    const int i = t / PF_CAND_PER_BLOB;
    const unsigned sp = span[i];
    const unsigned c = (sp & 255u) + (unsigned)(t - i * PF_CAND_PER_BLOB);
    if (c <= (sp >> 8)) {
      int rank = 0;
      for (int q = 0; q < i; ++q) { const unsigned sq = span[q]; rank += ((sq & 255u) <= c && c <= (sq >> 8)) ? 1 : 0; }
      cand[bt.start[c] + rank] = blobs[i];
    }
  }
  __syncthreads();
}

// nearest blob of a projected LED among the candidates of its column; m = +inf, am = 0 if there is none.
// The first record is tested without a branch (cand[0], the sentinel, for an empty column); the loop over further
// candidates is kept rolled: it runs zero times for most LEDs, and an unrolled remainder cascade costs more than it saves.
__device__ __forceinline__ void pf_nearest_one(const pf_blobtab& bt, const pf_blob* cand, float u, float v, float& m, int& am)
{
  const unsigned d = bt.desc[pf_col(bt, u)];
  const pf_blob* q = cand + (d & 0xFFFFu);
  const float4 z0 = *reinterpret_cast<const float4*>(q);
  const float dx0 = z0.x - u, dy0 = z0.y - v;
  float best = fmaf(dy0, dy0, dx0 * dx0);
  int bi = __float_as_int(z0.z);                            // a NaN distance (non-finite projection) is never accepted downstream
  if (d >= (2u << 16)) {
    const pf_blob* qe = q + (d >> 16);
#pragma unroll 1
    for (++q; q < qe; ++q) {
      const float4 z = *reinterpret_cast<const float4*>(q);
      const float dx = z.x - u, dy = z.y - v;
      const float d2 = fmaf(dy, dy, dx * dx);
      if (d2 < best) { best = d2; bi = __float_as_int(z.z); }
    }
  }
  m = best;
  am = bi;
}

// ---------------------------------------------------------------------------------------
// shared-memory frame context (uniform per CTA)
// ---------------------------------------------------------------------------------------
struct pf_ctx {
  pf_frame_const fc;
  float obj[PF_MAX_LED * 3];
  double objd[PF_MAX_LED * 3];
  int src_buf, src_mode, pad[2];
  pf_blobtab bt;
};

// ---------------------------------------------------------------------------------------
// Stage A: fused propagate -> project -> score over this warp's contiguous particles for
// iteration `it` (:543-605).  Leaves the warp's sign-agnostic prefix aggregate in warp_agg[warp].
// ---------------------------------------------------------------------------------------
template <int L, int V>
__device__ __forceinline__ pf_red pf_stage_score(const pf_kparams& p, const pf_ctx& cx, const pf_blob* blobs,
                                                 const uint32_t* tab, int trk, int it, pf_sm3* warp_agg)
{
  constexpr bool SH = V == 1, DIST = V == 2;   // kernel variants: 1 sharded stages, 2 plumb_bob projection
  const int N = p.N;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int S = p.pstride;
  const size_t tb = (size_t)trk * 3 * (size_t)S;
  // no __restrict__/const on buffers that other stages of this launch write
  const float4* src = (cx.src_buf == 0 ? p.pose0 : p.pose1) + tb;
  float4* dst = (cx.src_buf == 0 ? p.pose1 : p.pose0) + tb;
  float* w = p.w + (size_t)trk * N;
  // sharded filter: the rebalance delivers lead[] (slot -> local slot that holds its row), idx[] is the stage-B output
  const int* idx = (SH && p.lead) ? p.lead : p.idx + (size_t)trk * N;
  const pf_blob* cand = blobs + p.D_max;

  const bool use_cam = cx.fc.isi > 1;                       // :555-560
  const bool use_pred = use_cam && (it % 10) != 0;
  const bool cam_mul = use_cam && !(cx.fc.flags & PF_FC_CAM_IDENTITY);
  const int ang_mode = (cx.fc.flags & PF_FC_TINY_ANG) ? 2 : ((cx.fc.flags & PF_FC_SMALL_ANG) ? 1 : 0);
  const bool kstd = p.kstd != 0;
  const int steps = cx.fc.D < L ? cx.fc.D : L;              // :2409
  const bool gather = cx.src_mode == 1;
  const float scale = 1.0f + 0.025f * (float)(it / 10);    // :563 (integer division)

  const int wchunk = p.cta_chunk / (int)(blockDim.x >> 5); // multiple of 32
  const long long w0l = (long long)blockIdx.x * p.cta_chunk + (long long)warp * wchunk;
  const int w0 = w0l < (long long)N ? (int)w0l : N;
  const int w1 = (w0l + wchunk) < (long long)N ? (int)(w0l + wchunk) : N;
  const int goff = SH ? p.goff : 0;

  // RNG: particle n >= 2 starts at engine step 12*((N-2)*it + n-2).  Each lane keeps the
  // state of "its" particle of the current step and advances it by 32 particles = 384 engine steps
  // with one modular multiplication (exponents mod the group order, so n < 2 is harmless).
  uint32_t xs;
  {
    const uint64_t e0 = (12ull * (uint64_t)((SH ? p.N_global : N) - 2) * (uint64_t)it) % PF_RNG_ORDER;
    const uint64_t ee = (e0 + (uint64_t)PF_RNG_ORDER - 24ull + 12ull * (uint64_t)(goff + w0 + lane)) % PF_RNG_ORDER;
    xs = pf_rng_skip_e(cx.fc.seed_state, (uint32_t)ee, tab);
  }

  // per-thread running (max, first index) / (min, first index) as plain floats -- a thread visits its particles in
  // ascending order, so strict comparisons keep the first -- packed into the order-preserving keys once, at the end
  float bmax = -INFINITY, bmin = INFINITY;
  int imax = -1, imin = -1;
  long long tsum = 0;
  int s_next = (gather && w0 + lane < w1) ? idx[w0 + lane] : 0;   // Particle_index of the first step's slot (:678-681)
  for (int n = w0 + lane; n < w1; n += 32) {                // full steps except possibly the last: no warp-level code inside
    const uint32_t x = xs;
    xs = pf_modmul(xs, PF_RNG_A384);
    const int s = gather ? s_next : n;
    if (gather && n + 32 < w1) s_next = idx[n + 32];      // one step ahead: shortens the idx -> rows dependent chain
    pf_pose T;
    const int ng = goff + n;                               // global particle index (sharded filter: goff > 0)
    if (ng >= 2) {
      pf_pose B = pf_load_pose(src, S, s);
      if (cam_mul) { pf_pose Cm; for (int i = 0; i < 12; ++i) Cm.p[i] = cx.fc.C[i]; B = pf_affine_mul(Cm, B); }
      if (use_pred) { pf_pose Pm; for (int i = 0; i < 12; ++i) Pm.p[i] = cx.fc.P[i]; B = pf_affine_mul(B, Pm); }
      T = pf_propagate(B, x, cx.fc.noise, scale, ang_mode);
    } else {
      const float* t0 = ng == 0 ? cx.fc.Tcur : cx.fc.Tpred;   // :547, :551
#pragma unroll
      for (int i = 0; i < 12; ++i) T.p[i] = t0[i];
    }
    pf_store_pose(dst, S, n, T);
    // camera coordinates of every LED first, then one (uniform) branch on the shape of K for all of them
    float xc[L], yc[L], zc[L];
#pragma unroll
    for (int j = 0; j < L; ++j) pf_to_camera(T, cx.obj[3 * j], cx.obj[3 * j + 1], cx.obj[3 * j + 2], xc[j], yc[j], zc[j]);
    float u[L], v[L];
    if (kstd) {
#pragma unroll
      for (int j = 0; j < L; ++j) pf_to_pixel(p.K9f, true, xc[j], yc[j], zc[j], u[j], v[j]);
    } else {
#pragma unroll
      for (int j = 0; j < L; ++j) pf_to_pixel(p.K9f, false, xc[j], yc[j], zc[j], u[j], v[j]);
    }
    // the NEXT step's source rows on their way into L1 while this step searches its blobs: the rows hang on idx[n + 32]
    // (fetched at the top of this step), and the first use of a gathered row is the stage's largest single stall (9 % of
    // the samples).  Same-box A/B (PFMPE_PREFETCH=0 switches it off): 2^26 particles 2.79 -> 2.71 ms per frame, 1e6
    // particles with 205 blobs 88.9 -> 88.4 us, the batch shape unchanged (0.7256 -> 0.7250 ms).
    if (p.prefetch && n + 32 < w1) pf_prefetch_pose(src, S, gather ? s_next : n + 32);
    float m[L];
    int am[L];
#pragma unroll
    for (int j = 0; j < L; ++j) {
      if (DIST) pf_distort(p.K9f, p.dist, u[j], v[j]);     // row f3: score against the raw (distorted) blob centres
      pf_nearest_one(cx.bt, cand, u[j], v[j], m[j], am[j]);
    }
    const float wn = pf_weight_unordered<L>(m, am, steps, p.score);
    w[n] = wn;
    if (wn > bmax) { bmax = wn; imax = ng; }
    if (wn < bmin) { bmin = wn; imin = ng; }
    tsum += pf_w_to_fx(wn);
  }
  pf_red acc = pf_red_identity();
  acc.sum = tsum;
  if (imax >= 0) {
    acc.maxkey = ((unsigned long long)pf_ord(bmax) << 32) | (unsigned)(~(unsigned)imax);
    acc.minkey_inv = ~(((unsigned long long)pf_ord(bmin) << 32) | (unsigned)imin);
  }
  // The warp's sign-agnostic prefix aggregate (sum, running max, running min of the prefix of its particles in order).
  // Without a negative weight the prefix is non-decreasing: max = the total, min = the first element -- no per-step
  // reductions.  A warp that did see one (penalties can outweigh the gains, :2423-2432) re-reads its weights and runs
  // the ordered scan.
  pf_sm3 carry = pf_sm3_identity();
  __syncwarp();
  if (!__any_sync(PF_FULL, bmin < 0.0f)) {                 // the thread's minimum weight tells whether it saw a negative one
    long long tot = tsum;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(PF_FULL, tot, o);
    if (w0 < w1 && lane == 0) { carry.s = tot; carry.mx = tot; carry.mn = pf_w_to_fx(w[w0]); }   // lane 0 wrote w[w0] itself; only its carry is used
  } else {
#pragma unroll 1
    for (int base = w0; base < w1; base += 32) {
      const int n = base + lane;
      pf_sm3 a;
      if (n < w1) { const long long e = pf_w_to_fx(w[n]); a.s = e; a.mx = e; a.mn = e; } else a = pf_sm3_identity();
#pragma unroll 1
      for (int o = 1; o < 32; o <<= 1) {
        pf_sm3 t;
        t.s = __shfl_down_sync(PF_FULL, a.s, o);
        t.mx = __shfl_down_sync(PF_FULL, a.mx, o);
        t.mn = __shfl_down_sync(PF_FULL, a.mn, o);
        if (lane + o < 32) a = pf_sm3_combine(a, t);
      }
      a.s = pf_shfl_ll(a.s, 0); a.mx = pf_shfl_ll(a.mx, 0); a.mn = pf_shfl_ll(a.mn, 0);
      carry = pf_sm3_combine(carry, a);
    }
  }
  if (lane == 0) {
    warp_agg[warp] = carry;
    if ((SH && p.mode == 3)) {   // sharded filter: the resampling stage is a separate launch
      pf_cta_agg* g = p.warp_agg_g + (size_t)blockIdx.x * PF_MAX_WARPS + warp;
      g->s = carry.s; g->mx = carry.mx; g->mn = carry.mn;
    }
  }
  return acc;
}

// Prefix bookkeeping between the stages (after barrier 1), done by warp 0 and handed to the other
// warps through shared memory: scr[w] = (exclusive combination of the aggregates of CTAs 0..b-1)
// (+) (aggregates of warps 0..w-1 of this CTA) = everything before warp w; scr[PF_MAX_WARPS] = the
// combination of the whole filter.  Ends with a __syncthreads.
template <bool SH>
__device__ __noinline__ void pf_cta_prefix(const pf_kparams& p, const pf_sm3* warp_agg, pf_sm3* scr, const pf_xch* xlast)
{
  const int nwarps = blockDim.x >> 5;
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    pf_sm3 excl = pf_sm3_identity(), total = pf_sm3_identity();
    if (p.coop == 2) {                                    // the pair's aggregates of the last score pass (xlast[0], xlast[1])
      if (blockIdx.x == 1) excl = xlast[0].agg;
      total = pf_sm3_combine(xlast[0].agg, xlast[1].agg);
    } else if (p.coop) {
      // lane l owns the contiguous entries [l*seg, (l+1)*seg); loads are issued 4 at a time before use
      const int b = blockIdx.x, G = gridDim.x;
      const long long* g = reinterpret_cast<const long long*>(p.cta_agg);
      const int seg = (G + 31) / 32;
      const int i1 = (lane + 1) * seg < G ? (lane + 1) * seg : G;
      pf_sm3 v = pf_sm3_identity(), a = pf_sm3_identity();
#pragma unroll 1
      for (int i0 = lane * seg; i0 < i1; i0 += 4) {
        pf_sm3 t[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int i = i0 + q;
          if (i < i1) { t[q].s = __ldcg(g + 4 * i); t[q].mx = __ldcg(g + 4 * i + 1); t[q].mn = __ldcg(g + 4 * i + 2); }
          else t[q] = pf_sm3_identity();
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (i0 + q < b) v = pf_sm3_combine(v, t[q]);
          a = pf_sm3_combine(a, t[q]);
        }
      }
#pragma unroll 1
      for (int o = 1; o < 32; o <<= 1) {
        pf_sm3 t, u;
        t.s = __shfl_down_sync(PF_FULL, v.s, o); t.mx = __shfl_down_sync(PF_FULL, v.mx, o); t.mn = __shfl_down_sync(PF_FULL, v.mn, o);
        u.s = __shfl_down_sync(PF_FULL, a.s, o); u.mx = __shfl_down_sync(PF_FULL, a.mx, o); u.mn = __shfl_down_sync(PF_FULL, a.mn, o);
        if (lane + o < 32) { v = pf_sm3_combine(v, t); a = pf_sm3_combine(a, u); }
      }
      excl.s = pf_shfl_ll(v.s, 0); excl.mx = pf_shfl_ll(v.mx, 0); excl.mn = pf_shfl_ll(v.mn, 0);
      total.s = pf_shfl_ll(a.s, 0); total.mx = pf_shfl_ll(a.mx, 0); total.mn = pf_shfl_ll(a.mn, 0);
    }
    // inclusive ordered scan of this CTA's warp aggregates (nwarps <= 32: one per lane)
    pf_sm3 w = lane < nwarps ? warp_agg[lane] : pf_sm3_identity();
#pragma unroll 1
    for (int o = 1; o < 32; o <<= 1) {
      pf_sm3 t;
      t.s = __shfl_up_sync(PF_FULL, w.s, o); t.mx = __shfl_up_sync(PF_FULL, w.mx, o); t.mn = __shfl_up_sync(PF_FULL, w.mn, o);
      if (lane >= o) w = pf_sm3_combine(t, w);
    }
    pf_sm3 we;                                            // exclusive: aggregates of the warps below `lane`
    we.s = __shfl_up_sync(PF_FULL, w.s, 1); we.mx = __shfl_up_sync(PF_FULL, w.mx, 1); we.mn = __shfl_up_sync(PF_FULL, w.mn, 1);
    if (lane == 0) we = pf_sm3_identity();
    if ((SH && p.mode == 4)) {   // sharded filter: everything on the lower ranks comes first
      pf_sm3 re; re.s = p.rank_excl[0]; re.mx = p.rank_excl[1]; re.mn = p.rank_excl[2];
      excl = pf_sm3_combine(re, excl);
    }
    if (lane < nwarps) scr[lane] = pf_sm3_combine(excl, we);
    if (p.coop == 0) {                                    // one CTA: the total is the combination of its warps
      total.s = pf_shfl_ll(w.s, 31); total.mx = pf_shfl_ll(w.mx, 31); total.mn = pf_shfl_ll(w.mn, 31);
    }
    if ((SH && p.mode == 4)) { total.s = p.rank_total[0]; total.mx = p.rank_total[1]; total.mn = p.rank_total[2]; }
    if (lane == 0) scr[PF_MAX_WARPS] = total;
  }
  __syncthreads();
}

// one out-of-line copy of K(m) (two FP64 divisions inside): called twice per resampling stage
__device__ __noinline__ int pf_count_leq_dev(long long m, const pf_resample_cfg& rc, const uint32_t* tab)
{
  return pf_count_leq(m, rc, tab);
}

// ---------------------------------------------------------------------------------------
// Stage B: normalise + prefix + stratified resampling in source-centric form (:627-629, :668-682).
// The prefix runs on sg*Q.30 weights with the (sum, running max) operator; for every source i the
// cumulative offspring count kend[i] = K(M_i) is computed, count_i = kend[i] - kend[i-1], and the
// most-resampled particle (:685-686) falls out of the same pass.  Each warp walks its contiguous
// particles with a register carry whose start value is (exclusive CTA prefix) (+) (aggregates of
// the lower warps of this CTA), both from stage A.
// Every source then writes its own index into the output slots it owns, idx[kend[i-1] .. kend[i])
// (= Particle_index of :678-681), so that the next score pass reads its source with one coalesced
// load instead of searching.  The last source with offspring also owns the slots >= K_total
// ("not found keeps the previous index", Q7).  The 32 sources of a warp step own one contiguous
// slot range: the warp fills it with coalesced stores, each slot finding its source lane by a
// 5-step shuffle search over the lanes' cumulative counts; a source with more than PF_FILL_WARP
// slots is handed to the whole CTA after the walk (batch mode).  In grid mode a step that owns more
// than PF_FILL_STEP slots is listed instead and filled by the whole grid after the stage's barrier
// (pf_grid_fill).
// ---------------------------------------------------------------------------------------
#define PF_FILL_WARP 4096
#define PF_FILL_LIST 32
#define PF_FILL_STEP 256     /* grid mode: a step that owns more slots than this is listed for pf_grid_fill */
#define PF_FILL_SPLIT 8192   /* ... in pieces of at most this many slots */

struct pf_fill_list {
  int n;
  int start[PF_FILL_LIST], end[PF_FILL_LIST], src[PF_FILL_LIST];
};

template <bool SH>
__device__ __forceinline__ pf_red pf_stage_resample(const pf_kparams& p, const pf_ctx& cx, const uint32_t* tab, int trk,
                                                    long long S_fx, int iters_total, const pf_sm3* scr, pf_fill_list* fl)
{
  const int N = p.N;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float* w = p.w + (size_t)trk * N;
  int* kend = p.kend + (size_t)trk * N;
  int* idx = p.idx + (size_t)trk * N;
  const long long sg = S_fx < 0 ? -1ll : 1ll;
  const int Ng = SH ? p.N_global : N;
  const uint32_t x_base = pf_rng_skip(cx.fc.seed_state, 12ull * (uint64_t)(Ng - 2) * (uint64_t)iters_total, tab);
  const pf_resample_cfg rc = pf_make_resample_cfg(x_base, Ng, S_fx);

  const int wchunk = p.cta_chunk / (int)(blockDim.x >> 5);
  const long long w0l = (long long)blockIdx.x * p.cta_chunk + (long long)warp * wchunk;
  const int w0 = w0l < (long long)N ? (int)w0l : N;
  const int w1 = (w0l + wchunk) < (long long)N ? (int)(w0l + wchunk) : N;

  pf_red acc = pf_red_identity();
  if (w0 < w1) {
    pf_sm carry;
    pf_sm3_select(scr[warp], sg, carry.s, carry.m);
    // K(M) of everything before this warp and of the whole filter: one evaluation, two half-warps
    int k_carry, ktot, kbase = 0;     // kbase (sharded filter): first global slot owned by this rank's sources
    {
      long long ts_, tm_;
      pf_sm3_select(scr[PF_MAX_WARPS], sg, ts_, tm_);
      if (SH) {          // three values in one evaluation: lanes 0-10, 11-21, 22-31
        long long rs_ = 0, rm_ = PF_SM_NEG;
        if (p.mode == 4) {
          pf_sm3 re; re.s = p.rank_excl[0]; re.mx = p.rank_excl[1]; re.mn = p.rank_excl[2];
          pf_sm3_select(re, sg, rs_, rm_);
        }
        const int kk = pf_count_leq_dev(lane < 11 ? carry.m : (lane < 22 ? tm_ : rm_), rc, tab);
        k_carry = __shfl_sync(PF_FULL, kk, 0);
        ktot = __shfl_sync(PF_FULL, kk, 11);
        if (p.mode == 4) {
          kbase = __shfl_sync(PF_FULL, kk, 22);
          kbase = ktot == 0 ? p.goff : (kbase == ktot ? Ng : kbase);
          if (blockIdx.x == 0 && threadIdx.x == 0) p.shard_partial[1] = kbase;
        }
      } else {
        const int kk = pf_count_leq_dev(lane < 16 ? carry.m : tm_, rc, tab);
        k_carry = __shfl_sync(PF_FULL, kk, 0);
        ktot = __shfl_sync(PF_FULL, kk, 16);
      }
    }
    pf_stamp(p, 13);

    float w_next = w0 + lane < w1 ? w[w0 + lane] : 0.0f;
    for (int base = w0; base < w1; base += 32) {
      const int n = base + lane;
      const bool valid = n < w1;
      const float w_cur = w_next;
      if (n + 32 < w1) w_next = w[n + 32];                // one step ahead of the dependent chain scan -> K(M) -> fill
      const long long e0 = valid ? pf_w_to_fx(w_cur) : 0ll, e = sg < 0 ? -e0 : e0;
      long long M, total;
      if (!__any_sync(PF_FULL, e < 0)) {
        long long ps = e;                                 // inclusive prefix sum inside the step
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const long long t = __shfl_up_sync(PF_FULL, ps, o);
          if (lane >= o) ps += t;
        }
        const long long pfx = carry.s + ps;
        M = carry.m > pfx ? carry.m : pfx;
        total = pf_shfl_ll(ps, 31);
      } else {
        pf_sm a;
        if (valid) { a.s = e; a.m = e; } else a = pf_sm_identity();
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          pf_sm t;
          t.s = __shfl_up_sync(PF_FULL, a.s, o);
          t.m = __shfl_up_sync(PF_FULL, a.m, o);
          if (lane >= o) a = pf_sm_combine(t, a);
        }
        const pf_sm pre = pf_sm_combine(carry, a);
        M = pre.m;
        total = pf_shfl_ll(a.s, 31);
      }
      const int ke = pf_count_leq(M, rc, tab);
      int k_prev = __shfl_up_sync(PF_FULL, ke, 1);
      if (lane == 0) k_prev = k_carry;
      if (valid) {
        kend[n] = ke;
        const unsigned cnt = (unsigned)(ke - k_prev);
        const unsigned long long key = ((unsigned long long)cnt << 32) | (unsigned)(~(unsigned)((SH ? p.goff : 0) + n));
        acc.cntkey = key > acc.cntkey ? key : acc.cntkey;
      }
      // ---- slot fill: the sources of this step own the contiguous slots [E(k_carry), E(ke of lane 31)),
      // E(k) = N if k == K_total (the last source with offspring also owns the not-found tail, Q7)
      if (ktot == 0) {
        if (valid) idx[n] = n;                            // nothing was ever found: every slot keeps its own particle (local)
        if ((SH && p.mode == 4) && valid && n == N - 1) p.shard_partial[2] = (long long)p.goff + N;
      } else {
        const int Ecur = (ke == ktot ? Ng : ke) - kbase;
        const int Eprev = (k_prev == ktot ? Ng : k_prev) - kbase;
        const int own = Ecur - Eprev;                     // invalid lanes repeat the last prefix: 0
        if ((SH && p.mode == 4) && valid && n == N - 1) p.shard_partial[2] = (long long)Ecur + kbase;   // end of this rank's slots
        const int E0 = __shfl_sync(PF_FULL, Eprev, 0), R = __shfl_sync(PF_FULL, Ecur, 31) - E0;
        bool listed = false;
        if (p.coop == 1 && R > PF_FILL_STEP) {
          // heavy step (grid mode): its slots are filled after the stage's barrier by whichever warp of
          // the grid picks the entry up (pf_grid_fill); pieces of at most PF_FILL_SPLIT slots
          const int parts = (R + PF_FILL_SPLIT - 1) / PF_FILL_SPLIT;
          int q = 0;
          if (lane == 0) {
            q = atomicAdd(&p.gfill->n, parts);
            if (q + parts > p.gfill_cap) { atomicSub(&p.gfill->n, parts); q = -1; }
            else { p.gfill->ktot = ktot; p.gfill->kbase = kbase; acc.sum += parts; }   // the count rides on the stage's reduction
          }
          q = __shfl_sync(PF_FULL, q, 0);
          if (q >= 0) {
            for (int j = lane; j < parts; j += 32) {
              const int off = j * PF_FILL_SPLIT;
              p.gfill->ent[q + j] = make_int4(base, E0 + off, R - off < PF_FILL_SPLIT ? R - off : PF_FILL_SPLIT, 0);
            }
            listed = true;
          }
        }
        const unsigned bigm = p.coop == 1 ? 0u : __ballot_sync(PF_FULL, own > PF_FILL_WARP);
        const int maxown = listed ? 0 : __reduce_max_sync(PF_FULL, own);
        if (listed) {
        } else if (bigm == 0u && maxown <= 6 * ((R + 31) >> 5)) {
          // balanced step (the usual case: offspring counts of a few): every source writes its own slots, one per
          // round; the lanes' slot ranges are adjacent, so a round's stores fall into a window of at most R slots
          int* q = idx + Eprev;
          const int lim = SH ? min(own, p.idx_cap - Eprev) : own;
#pragma unroll
          for (int t = 0; t < 4; ++t) if (t < lim) q[t] = n;          // predicated stores, no loop
          if (maxown > 4) {
#pragma unroll 1
            for (int t = 4; t < maxown; ++t) if (t < lim) q[t] = n;   // warp-uniform trip count
          }
        } else if (bigm == 0u) {
          for (int q0 = 0; q0 < R; q0 += 32) {            // warp-uniform trip count: every lane serves the shuffles
            const int k = E0 + q0 + lane;
            int lo = 0, hi = 31;                          // smallest lane whose Ecur > k
#pragma unroll
            for (int t = 0; t < 5; ++t) {
              const int mid = (lo + hi) >> 1;
              const int v = __shfl_sync(PF_FULL, Ecur, mid);
              if (v > k) hi = mid; else lo = mid + 1;
            }
            if (q0 + lane < R && (!SH || k < p.idx_cap)) idx[k] = base + lo;
          }
        } else {
          // one CTA per filter (batch mode): a source with more than PF_FILL_WARP slots is handed to the
          // whole CTA (list, after the walk); every other source of this step -- and list overflow -- is
          // filled by the whole warp, one source at a time (warp-uniform loop: the mask is a ballot)
          const bool isbig = (bigm >> lane) & 1u;
          bool handed = false;
          if (isbig) {
            const int slot = atomicAdd(&fl->n, 1);
            if (slot < PF_FILL_LIST) { fl->start[slot] = Eprev; fl->end[slot] = Ecur; fl->src[slot] = n; handed = true; }
          }
          unsigned todo = __ballot_sync(PF_FULL, own > 0 && !handed);
          while (todo) {
            const int l = __ffs(todo) - 1;
            todo &= todo - 1u;
            const int s0 = __shfl_sync(PF_FULL, Eprev, l), s1 = __shfl_sync(PF_FULL, Ecur, l);
            for (int k = s0 + lane; k < s1; k += 32)
              if (!SH || k < p.idx_cap) idx[k] = base + l;
          }
        }
      }
      carry.s += total;
      carry.m = pf_shfl_ll(M, 31);          // invalid lanes repeat the last valid prefix
      k_carry = __shfl_sync(PF_FULL, ke, 31);
    }
  }
  pf_stamp(p, 14);
  __syncthreads();
  {
    const int nl = fl->n < PF_FILL_LIST ? fl->n : PF_FILL_LIST;
    for (int q = 0; q < nl; ++q) {
      const int s0 = fl->start[q], s1 = fl->end[q], src = fl->src[q];
      for (int k = s0 + (int)threadIdx.x; k < s1 && (!SH || k < p.idx_cap); k += (int)blockDim.x) idx[k] = src;
    }
  }
  return acc;
}

// Grid-wide part of the slot fill (grid mode).  A source-centric fill is only as balanced as the
// offspring counts: after a few frames the heavy particles of a large filter are neighbours (they
// descend from the same few particles), so a handful of warps would write most of idx[].  Heavy
// steps are therefore listed during the walk and filled here, after the barrier that ends the
// resampling stage, by all warps of the grid round-robin, with the same shuffle search over the 32
// cumulative counts (re-read from kend[]).  The list length comes out of that stage's grid
// reduction (pf_red.sum), so every CTA sees the same value and takes the same decision about the
// extra barrier that orders this fill before the next score pass reads idx[].
__device__ __noinline__ void pf_grid_fill(const pf_kparams& p, int trk, pf_bar& bar, int n, int Ng, bool check_cap)
{
  const int N = p.N;
  const int ktot = __ldcg(&p.gfill->ktot), kbase = __ldcg(&p.gfill->kbase);
  const int* kend = p.kend + (size_t)trk * N;
  int* idx = p.idx + (size_t)trk * N;
  const int lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int tw = (int)gridDim.x * nw;
#pragma unroll 1
  for (int q = (int)blockIdx.x + (int)(threadIdx.x >> 5) * (int)gridDim.x; q < n; q += tw) {   // neighbouring entries -> different SMs
    const int4 e = __ldcg(&p.gfill->ent[q]);
    const int nn = e.x + lane < N ? e.x + lane : N - 1;
    const int ke = __ldcg(kend + nn);
    const int Ecur = (ke == ktot ? Ng : ke) - kbase;
#pragma unroll 1
    for (int q0 = 0; q0 < e.z; q0 += 32) {
      const int k = e.y + q0 + lane;
      int lo = 0, hi = 31;
#pragma unroll
      for (int t = 0; t < 5; ++t) {
        const int mid = (lo + hi) >> 1;
        const int v = __shfl_sync(PF_FULL, Ecur, mid);
        if (v > k) hi = mid; else lo = mid + 1;
      }
      if (q0 + lane < e.z && (!check_cap || k < p.idx_cap)) idx[k] = e.x + lo;
    }
  }
  pf_grid_barrier(p, bar);
  if (blockIdx.x == 0 && threadIdx.x == 0) p.gfill->n = 0;
}

// ---------------------------------------------------------------------------------------
// Stage C: tail of one frame, executed by warp 0 of every CTA of the tracker (redundantly, so
// that no third grid barrier is needed: every CTA derives the next frame's constants itself);
// only CTA 0 writes to global memory.  Winner, correspondences, result record, bookkeeping; in
// track mode also Gauss-Newton + updatePose + next frame's prediction
// (:685-703, :707-719, :2011-2035, :995-1010).
// ---------------------------------------------------------------------------------------
template <int L, int V>
__device__ __noinline__ void pf_frame_tail(const pf_kparams& p, pf_ctx& cx, pf_tracker_state& st, pf_tail_smem& ts,
                                           const pf_blob* blobs, int trk, int frame, int iters, int best_it,
                                           float highest, const pf_red& kept, bool accept, bool resampled,
                                           const pf_red& cnt)
{
  constexpr bool SH = V == 1, DIST = V == 2;
  const int N = p.N;
  const int lane = threadIdx.x & 31;
  const bool writer = blockIdx.x == 0 && lane == 0;
  const bool seq_mode = p.mode == 1 || p.mode == 2;
  pf_frame_result* res = p.result + (seq_mode ? (size_t)trk * p.total_frames + frame : (size_t)trk);
  const int dst_buf = (SH && p.mode == 5) ? st.t_buf : 1 - cx.src_buf;   // sharded winner stage: a launch of its own, after the resampling may have made T the source
  const float4* T = (dst_buf == 0 ? p.pose0 : p.pose1) + (size_t)trk * 3 * (size_t)p.pstride;

  const int arg_max = (int)(~(unsigned)(kept.maxkey & 0xFFFFFFFFull));
  const int arg_min = (int)(unsigned)((~kept.minkey_inv) & 0xFFFFFFFFull);
  int q;
  if (resampled) q = (int)(~(unsigned)(cnt.cntkey & 0xFFFFFFFFull));   // :686 first max of counterMeas
  else q = kept.sum < 0 ? arg_min : arg_max;                            // :714 on the normalised weights (Q6)
  if ((SH && p.mode == 5)) q = p.q_local;                                       // sharded: chosen across the ranks by the host

  const pf_pose Tw = pf_load_pose(T, p.pstride, q);
  unsigned pairs[2 * L];
  int npairs = 0;
#pragma unroll
  for (int j = 0; j < 2 * L; ++j) pairs[j] = 0u;
  (void)pf_score_warp<L>(blobs, cx.fc.D, p.K9f, p.kstd != 0, DIST ? p.dist : nullptr, Tw, cx.obj, p.score, pairs, &npairs, lane);
  if (writer) {
    res->status = accept ? 1 : 4;
    res->iterations = iters;
    res->best_iteration = best_it;
    res->resampled = resampled ? 1 : 0;
    res->highest_prob = highest;
    res->most_likely_particle_idx = accept ? (best_it >= 0 ? arg_max : -1) : q;
    res->most_likely_pose_idx = resampled ? q : -1;
    res->max_count = resampled ? (unsigned)(cnt.cntkey >> 32) : 0u;
    res->sum_fx = kept.sum;
#pragma unroll
    for (int i = 0; i < 12; ++i) res->winner_pose[i] = Tw.p[i];
#pragma unroll
    for (int i = 0; i < 2 * PF_MAX_LED; ++i) res->pairs[i] = 0u;
#pragma unroll
    for (int i = 0; i < 2 * L; ++i) res->pairs[i] = i < 2 * npairs ? pairs[i] : 0u;
    res->npairs = npairs;
    res->gn_iters = -1;
    res->flag_fail = accept ? 1.0f : 4.0f;
    if (p.done_seq) {     // per-frame latency path: the host polls this word instead of waiting for the kernel to retire
      __threadfence_system();
      *reinterpret_cast<volatile int32_t*>(&res->pad) = p.done_seq;
    }
  }
  __syncwarp();
  pf_stamp(p, 7);

  if (p.mode == 1 && !p.split_tail) {
    // ---- device tail (track mode): optimiseAndUpdatePose + next predictPose -------------
    pf_aff cur, prev;
#pragma unroll
    for (int i = 0; i < 12; ++i) { cur.a[i] = st.cur[i]; prev.a[i] = st.prev[i]; }
    double t_cur = st.t_cur, t_prev = st.t_prev;
    const double t_pred = st.t_pred;
    uint32_t isi = st.isi;
    __syncwarp();
    if (resampled) {
      pf_aff pose = pf_aff_from_f32(Tw);                 // predicted_pose_ = PoseParticle[q]  (:687)
      const pf_aff before = pose;
      const double* img = p.blobs + ((size_t)trk * p.total_frames + frame) * (size_t)p.D_max * 2;
      const int gn = pf_gn_warp<L>(p.K9d, cx.objd, img, pairs, npairs, pose, ts, lane, p.gn_cap);   // cov -> ts.M
      pf_stamp(p, 10);
      if (isi < 2) ++isi;                                // :2027-2030
      prev = cur;                                        // updatePose :2013-2014
      cur = pose;
      if (t_pred - t_cur > 0.001 || t_pred < t_cur) { t_prev = t_cur; t_cur = t_pred; }   // :2016-2020
      float flag = 1.0f;
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j)
          if (fabs(before.a[4 * i + j] - pose.a[4 * i + j]) >= 0.3) flag = 1.5f;           // :693-701
      if (writer) { res->flag_fail = flag; res->gn_iters = gn; }
      for (int i = lane; i < 36; i += 32) {
        const double c = ts.M[i];
        st.cov[i] = c;
        if (blockIdx.x == 0) res->cov[i] = c;
      }
      if (lane == 0) {
#pragma unroll
        for (int i = 0; i < 12; ++i) st.pred[i] = pose.a[i];
      }
    } else {
      isi = 0;                                           // :710
      if (lane == 0) {
        st.failed = 1;
#pragma unroll
        for (int i = 0; i < 12; ++i) st.pred[i] = (double)Tw.p[i];   // :716
      }
    }
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 12; ++i) { st.cur[i] = cur.a[i]; st.prev[i] = prev.a[i]; }
      st.t_cur = t_cur; st.t_prev = t_prev; st.isi = isi;
    }
    if (writer) {
#pragma unroll
      for (int i = 0; i < 12; ++i) res->pose[i] = cur.a[i];
    }
    if (resampled && frame + 1 < p.total_frames) {
      const double t_next = p.times[frame + 1];
      pf_aff P, predn;
      pf_predict_pose_dev(prev, cur, t_prev, t_cur, t_next, P, predn);   // :234 -> :995-1010
      pf_stamp(p, 11);
      const size_t fi = (size_t)trk * p.total_frames + frame + 1;
      if (lane == 0) {
        st.t_pred = t_next;
#pragma unroll
        for (int i = 0; i < 12; ++i) st.pred[i] = predn.a[i];
        pf_make_frame_const(cx.fc, cur, predn, P, pf_aff_identity(), isi, t_next - t_cur, p.seeds[fi],
                            p.blob_count[fi], p.min_ang, p.max_ang, p.min_tr, p.max_tr);
      }
    }
    __syncwarp();
  }
  // CTA 0 mirrors the tracker state (and the staged constants of the next frame) to global memory
  // for the next launch; nothing reads them again during this one.
  if (blockIdx.x == 0) {
    const uint32_t* s32 = reinterpret_cast<const uint32_t*>(&st);
    uint32_t* g32 = reinterpret_cast<uint32_t*>(p.state + trk);
    for (int i = lane; i < (int)(sizeof(pf_tracker_state) / 4); i += 32) g32[i] = s32[i];
    if (p.mode == 1 && !p.split_tail) {
      const uint32_t* f32 = reinterpret_cast<const uint32_t*>(&cx.fc);
      uint32_t* h32 = reinterpret_cast<uint32_t*>(p.fconst + trk);
      for (int i = lane; i < (int)(sizeof(pf_frame_const) / 4); i += 32) h32[i] = f32[i];
    }
  }
}

// Sharded filter: CTA 0 delivers this rank's partial (nwords payload words + sequence number) to the
// mailbox of every rank -- its own included -- with peer stores; the sequence word is written last,
// after a system-scope fence, and the receivers poll for it (pf_shard_wait_kernel).  Replaces the
// NCCL all_gather + device->host copy between the stages.
__device__ __forceinline__ void pf_shard_publish(const pf_kparams& p, int nwords)
{
  if (!p.xmail_peers || blockIdx.x != 0) return;
  __syncthreads();                                   // the partial written by this CTA's thread 0
  if ((int)threadIdx.x < p.xworld) {
    long long* dst = p.xmail_peers[threadIdx.x] + ((size_t)p.xslot * p.xworld + p.xrank) * PF_XMAIL_WORDS;
    for (int i = 0; i < nwords; ++i) dst[i] = __ldcg(p.shard_partial + i);
    __threadfence_system();
    *reinterpret_cast<volatile long long*>(dst + PF_XMAIL_WORDS - 1) = p.xseq;
  }
}

// ---------------------------------------------------------------------------------------
// The persistent frame kernel.
// ---------------------------------------------------------------------------------------
// Variants (separate instantiations, so that the plain single-GPU kernel carries none of their code):
// V = 1: the stage-per-launch forms of the sharded filter (modes 3-5); V = 2: forward plumb_bob
// distortion in the projection (row f3).
// V = 3: the plain kernel again, compiled for the one-CTA-per-filter (batch) shape only: 256-thread CTAs, 4 per SM
// (64 registers instead of 80): the batch shape is bound by issue latency, and a third more resident warps buys
// 7-10 % on BASELINE configs[3] (4096 filters x 4096 particles: 1.43 -> 1.32 ms/step).
template <int L, int V>
__global__ void __launch_bounds__(V == 3 ? PF_BLOCK : PF_MAX_BLOCK, V == 3 ? 4 : 1) pf_frames_kernel(const __grid_constant__ pf_kparams p)
{
  constexpr bool SH = V == 1;
  extern __shared__ __align__(16) unsigned char pf_dyn_smem[];
  pf_blob* blobs = reinterpret_cast<pf_blob*>(pf_dyn_smem);
  __shared__ pf_ctx cx;
  __shared__ pf_tracker_state st;
  __shared__ pf_tail_smem ts;
  __shared__ pf_red red_smem[PF_MAX_WARPS + 1];   // [PF_MAX_WARPS]: grid result broadcast
  __shared__ pf_sm3 warp_agg[PF_MAX_WARPS];
  __shared__ pf_sm3 scr[PF_MAX_WARPS + 1];
  __shared__ pf_fill_list fill;
  __shared__ pf_xch xch[2][2];          // cluster-pair mode: [round parity][CTA of the pair]
  __shared__ uint32_t tab[PF_RNG_TAB_WORDS];   // radix-512 power tables of the RNG skip-ahead

  const int trk = p.tracker_base + blockIdx.y;
  const int N = p.N;
  // reduction-slot rotation and the barrier count continue across launches (accum[3]); every CTA
  // reads them before its first barrier, CTA 0 writes them back after its last one
  pf_bar bar;
  bar.round = p.coop == 1 ? (unsigned)(*(volatile unsigned long long*)&p.accum[3].maxkey) : 0u;
  bar.done = p.coop == 1 ? *(volatile unsigned long long*)&p.accum[3].minkey_inv : 0ull;

  // tracker state and the staged constants of the first frame: global -> shared (every CTA)
  {
    const uint32_t* g32 = reinterpret_cast<const uint32_t*>(p.state + trk);
    uint32_t* s32 = reinterpret_cast<uint32_t*>(&st);
    for (int i = threadIdx.x; i < (int)(sizeof(pf_tracker_state) / 4); i += blockDim.x) s32[i] = g32[i];
    const uint32_t* h32 = reinterpret_cast<const uint32_t*>(p.inline_in ? &p.fc_inline : p.fconst + trk);
    uint32_t* f32 = reinterpret_cast<uint32_t*>(&cx.fc);
    for (int i = threadIdx.x; i < (int)(sizeof(pf_frame_const) / 4); i += blockDim.x) f32[i] = h32[i];
    if (threadIdx.x < 3 * L) {
      cx.obj[threadIdx.x] = p.obj_f[(size_t)trk * PF_MAX_LED * 3 + threadIdx.x];
      cx.objd[threadIdx.x] = p.obj_d[(size_t)trk * PF_MAX_LED * 3 + threadIdx.x];
    }
    for (int i = threadIdx.x; i < PF_RNG_TAB_WORDS; i += blockDim.x) tab[i] = p.rng_tab[i];
  }

  for (int f = 0; f < p.n_frames; ++f) {
    const int frame = p.frame0 + f;
    pf_stamp(p, 0);
    __syncthreads();       // state / constants (initial load or the previous frame's tail) visible
    if (p.mode == 1 && st.failed) {   // track lost earlier: the host must re-initialise (:707-719)
      if (blockIdx.x == 0 && threadIdx.x == 0) p.result[(size_t)trk * p.total_frames + frame].status = 0;
      continue;
    }
    if (p.mode == 2) {     // replay: the frame's constants were logged by an earlier track-mode run
      const uint32_t* h32 = reinterpret_cast<const uint32_t*>(p.fc_log + (size_t)trk * p.total_frames + frame);
      uint32_t* f32 = reinterpret_cast<uint32_t*>(&cx.fc);
      for (int i = threadIdx.x; i < (int)(sizeof(pf_frame_const) / 4); i += blockDim.x) f32[i] = h32[i];
      __syncthreads();
    } else if (p.mode == 1 && p.fc_log && blockIdx.x == 0) {
      const uint32_t* f32 = reinterpret_cast<const uint32_t*>(&cx.fc);
      uint32_t* h32 = reinterpret_cast<uint32_t*>(p.fc_log + (size_t)trk * p.total_frames + frame);
      for (int i = threadIdx.x; i < (int)(sizeof(pf_frame_const) / 4); i += blockDim.x) h32[i] = f32[i];
    }
    if (threadIdx.x == 0) { cx.src_buf = st.src_buf; cx.src_mode = st.src_mode; }   // visible after the barriers of the staging
    const int D = cx.fc.D;
    {
      const bool seq_mode = p.mode == 1 || p.mode == 2;
      const double* bsrc = p.inline_in ? p.blobs_inline
                                       : p.blobs + ((size_t)trk * (seq_mode ? p.total_frames : 1) + (seq_mode ? frame : 0)) *
                                                       (size_t)p.D_max * 2;
      pf_stage_blobs(bsrc, D, p.score.tol_pf, cx.bt, blobs, p.D_max);
    }
    pf_stamp(p, 1);

    // ---- do { propagate, project, score } while (iter < 80 && max < L*min(5,numLED))  (:535-616)
    const float thr_exit = (float)(L * (D < 5 ? D : 5));
    float highest = 0.0f;
    int best_it = -1, it = 0;
    pf_red kept = pf_red_identity(), last = pf_red_identity();
    int iters = 0;
    bool regen = false;                 // extra pass that regenerates the kept iteration (one call site: one copy of the code)
    // Sharded single filter (modes 3-5): one stage per launch; the collectives across the GPUs and the
    // do/while decision sit between the launches (pf_monocular_pose_estimator_b200/sharded.py).
    while (!SH || p.mode < 4) {
      pf_red r = pf_stage_score<L, V>(p, cx, blobs, tab, trk, (SH && p.mode == 3) ? p.pass_it : (regen ? best_it : it), warp_agg);
      if (it == 0) pf_stamp(p, 2);
      r = pf_grid_reduce(r, red_smem, p, bar, warp_agg, xch);
      if (it == 0) pf_stamp(p, 3);
      if ((SH && p.mode == 3)) {   // this shard's (max, first argmax), (min, first argmin), sum and prefix aggregate
        pf_cta_prefix<SH>(p, warp_agg, scr, xch[(bar.round + 1u) & 1u]);
        if (blockIdx.x == 0 && threadIdx.x == 0) {
          p.state[trk].t_buf = 1 - cx.src_buf;        // where this pass left PoseParticle (getters, winner stage)
          p.shard_partial[0] = (long long)r.maxkey; p.shard_partial[1] = (long long)r.minkey_inv; p.shard_partial[2] = r.sum;
          p.shard_partial[3] = scr[PF_MAX_WARPS].s; p.shard_partial[4] = scr[PF_MAX_WARPS].mx; p.shard_partial[5] = scr[PF_MAX_WARPS].mn;
        }
        pf_shard_publish(p, 6);
        break;
      }
      if (regen) break;                                              // r == kept (same particles, exact sums)
      const float m = pf_unord((unsigned)(r.maxkey >> 32));
      last = r;
      ++it;
      if (m > highest) { highest = m; best_it = it - 1; kept = r; }   // :608-614 (strict >, from 0)
      if (it < PF_MAX_ITER && m < thr_exit) continue;                // :616
      iters = it;
      if (best_it < 0) { kept = last; break; }                       // :619: nothing to restore
      if (best_it == iters - 1) break;
      // the kept iteration is not the last one: regenerate its particles, weights and prefix
      // aggregates (exact: the RNG is addressed by (iteration, particle))
      regen = true;
    }

    if ((SH && p.mode == 3)) continue;
    if (SH && p.mode >= 4) {     // the host made the loop / accept decisions from the gathered partials
      kept.sum = p.S_fx; iters = p.iters_total; highest = INFINITY;
      if ((SH && p.mode == 4)) {   // prefix aggregates of the kept score pass (an earlier launch)
        const int nw = blockDim.x >> 5;
        if ((int)threadIdx.x < nw) {
          const pf_cta_agg* g = p.warp_agg_g + (size_t)blockIdx.x * PF_MAX_WARPS + threadIdx.x;
          pf_sm3 a; a.s = g->s; a.mx = g->mx; a.mn = g->mn;
          warp_agg[threadIdx.x] = a;
        }
        __syncthreads();
      }
    }

    // ---- accept test (:633) and resampling (:666-682)
    const bool accept = (kept.sum != 0) && (highest > (float)(L * (D < 3 ? D : 3)));
    const bool resampled = accept && cx.fc.isi > 0;
    pf_red cnt = pf_red_identity();
    if (resampled && !(SH && p.mode == 5)) {
      pf_stamp(p, 4);
      if (threadIdx.x == 0) fill.n = 0;                  // ordered before its use by the barrier in pf_cta_prefix
      pf_cta_prefix<SH>(p, warp_agg, scr, xch[(bar.round + 1u) & 1u]);
      pf_stamp(p, 12);
      cnt = pf_stage_resample<SH>(p, cx, tab, trk, kept.sum, iters, scr, &fill);
      pf_stamp(p, 5);
      cnt = pf_grid_reduce(cnt, red_smem, p, bar, nullptr, xch);
      if (p.coop == 1 && cnt.sum > 0) pf_grid_fill(p, trk, bar, (int)cnt.sum, SH ? p.N_global : N, SH);
      pf_stamp(p, 6);
    }

    if ((SH && p.mode == 4)) {     // sharded: the resampled set is rebalanced across the ranks into the source buffer
      if (blockIdx.x == 0 && threadIdx.x == 0) {
        p.shard_partial[0] = (long long)cnt.cntkey;
        pf_tracker_state* g = p.state + trk;
        // peer-memory rebalance: the propagated set itself becomes the source (rows that stay on their rank are not moved;
        // lead[] indexes the shard's own rows [0, N) and the receive area [N, 2N) behind them).  Row exchange: full rows
        // are written over the old source set, no indirection.
        g->t_buf = 1 - cx.src_buf; g->frames_done = st.frames_done + 1;
        if (p.lead_mode) { g->src_buf = 1 - cx.src_buf; g->src_mode = 1; } else g->src_mode = 0;
      }
      pf_shard_publish(p, 2);
      continue;
    }
    // particle-set bookkeeping for the next frame (shared memory copy of every CTA; CTA 0 mirrors it)
    if (threadIdx.x == 0 && !(SH && p.mode == 5)) {
      st.t_buf = 1 - cx.src_buf;
      if (resampled) { st.src_buf = 1 - cx.src_buf; st.src_mode = 1; }
      st.frames_done += 1;
    }
    __syncwarp();
    if (threadIdx.x < 32 && (p.mode == 1 || blockIdx.x == 0))
      pf_frame_tail<L, V>(p, cx, st, ts, blobs, trk, frame, iters, best_it, highest, kept, accept, resampled, cnt);
    if (SH && p.mode == 5 && p.xmail_peers) {   // winner pose + correspondences to every rank: 15 words
      if (blockIdx.x == 0 && threadIdx.x == 0) {
        const pf_frame_result* res = p.result + trk;
        const unsigned* wp = reinterpret_cast<const unsigned*>(res->winner_pose);
        for (int i = 0; i < 6; ++i) p.shard_partial[i] = (long long)(((unsigned long long)wp[2 * i + 1] << 32) | wp[2 * i]);
        for (int i = 0; i < PF_MAX_LED; ++i)
          p.shard_partial[6 + i] = (long long)(((unsigned long long)res->pairs[2 * i + 1] << 32) | res->pairs[2 * i]);
        p.shard_partial[6 + PF_MAX_LED] = res->npairs;
      }
      pf_shard_publish(p, 7 + PF_MAX_LED);
    }
    pf_stamp(p, 8);
    pf_stamp(p, 9);
  }
  if (p.coop == 1 && blockIdx.x == 0 && threadIdx.x == 0) {
    p.accum[3].maxkey = (unsigned long long)(bar.round % 3u);
    p.accum[3].minkey_inv = bar.done;
  }
}

// Track mode, one frame per launch, a batch of filters (p.split_tail): the FP64 tail of pf_frame_tail as a kernel of its
// own, ONE WARP PER FILTER instead of one warp of a 256-thread CTA whose other seven wait -- 4096 filters keep 4096 warps
// busy instead of 592 CTA slots, and the FP64 code (a third of the fused kernel's 190 KB) stays out of the instruction
// cache of the streaming stages.  Reads the frame's record (winner pose, correspondences, accept / resample flags) and
// the tracker state the frame kernel mirrored to global memory; same arithmetic as the fused form (pf_gn_warp,
// pf_predict_pose_dev), same results.
#define PF_TAIL_WARPS 4
template <int L>
__global__ void __launch_bounds__(32 * PF_TAIL_WARPS) pf_track_tail_kernel(const __grid_constant__ pf_kparams p, int frame)
{
  __shared__ pf_tail_smem tsm[PF_TAIL_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int trk = p.tracker_base + blockIdx.x * PF_TAIL_WARPS + warp;
  if (trk >= p.tracker_base + p.tail_count) return;
  pf_tail_smem& ts = tsm[warp];
  pf_tracker_state* st = p.state + trk;
  pf_frame_result* res = p.result + (size_t)trk * p.total_frames + frame;
  if (res->status == 0) return;                          // track lost earlier: the frame kernel skipped this filter
  const bool resampled = res->resampled != 0;
  pf_pose Tw;
#pragma unroll
  for (int i = 0; i < 12; ++i) Tw.p[i] = res->winner_pose[i];
  unsigned pairs[2 * L];
#pragma unroll
  for (int j = 0; j < 2 * L; ++j) pairs[j] = res->pairs[j];
  const int npairs = res->npairs;
  pf_aff cur, prev;
#pragma unroll
  for (int i = 0; i < 12; ++i) { cur.a[i] = st->cur[i]; prev.a[i] = st->prev[i]; }
  double t_cur = st->t_cur, t_prev = st->t_prev;
  const double t_pred = st->t_pred;
  uint32_t isi = st->isi;
  __syncwarp();
  if (resampled) {
    pf_aff pose = pf_aff_from_f32(Tw);                   // predicted_pose_ = PoseParticle[q]  (:687)
    const pf_aff before = pose;
    const double* img = p.blobs + ((size_t)trk * p.total_frames + frame) * (size_t)p.D_max * 2;
    const int gn = pf_gn_warp<L>(p.K9d, p.obj_d + (size_t)trk * PF_MAX_LED * 3, img, pairs, npairs, pose, ts, lane, p.gn_cap);   // cov -> ts.M
    if (isi < 2) ++isi;                                  // :2027-2030
    prev = cur;                                          // updatePose :2013-2014
    cur = pose;
    if (t_pred - t_cur > 0.001 || t_pred < t_cur) { t_prev = t_cur; t_cur = t_pred; }   // :2016-2020
    float flag = 1.0f;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j)
        if (fabs(before.a[4 * i + j] - pose.a[4 * i + j]) >= 0.3) flag = 1.5f;           // :693-701
    if (lane == 0) { res->flag_fail = flag; res->gn_iters = gn; }
    for (int i = lane; i < 36; i += 32) {
      const double c = ts.M[i];
      st->cov[i] = c;
      res->cov[i] = c;
    }
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 12; ++i) st->pred[i] = pose.a[i];
    }
  } else {
    isi = 0;                                             // :710
    if (lane == 0) {
      st->failed = 1;
#pragma unroll
      for (int i = 0; i < 12; ++i) st->pred[i] = (double)Tw.p[i];   // :716
    }
  }
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < 12; ++i) { st->cur[i] = cur.a[i]; st->prev[i] = prev.a[i]; res->pose[i] = cur.a[i]; }
    st->t_cur = t_cur; st->t_prev = t_prev; st->isi = isi;
  }
  if (resampled && frame + 1 < p.total_frames) {
    const double t_next = p.times[frame + 1];
    pf_aff P, predn;
    pf_predict_pose_dev(prev, cur, t_prev, t_cur, t_next, P, predn);   // :234 -> :995-1010
    const size_t fi = (size_t)trk * p.total_frames + frame + 1;
    if (lane == 0) {
      st->t_pred = t_next;
#pragma unroll
      for (int i = 0; i < 12; ++i) st->pred[i] = predn.a[i];
      pf_make_frame_const(p.fconst[trk], cur, predn, P, pf_aff_identity(), isi, t_next - t_cur, p.seeds[fi],
                          p.blob_count[fi], p.min_ang, p.max_ang, p.min_tr, p.max_tr);
    }
  }
}

// Materialise newPoseEstimation (the resampled set) for the getters / ROI: out[n] = T[idx[n]].
__global__ void __launch_bounds__(PF_BLOCK) pf_gather_kernel(const float4* T, const int* idx, int N, int S, int src_mode,
                                                             float4* out, int* idx_out)
{
  const int n = blockIdx.x * PF_BLOCK + threadIdx.x;
  if (n >= N) return;
  const int s = src_mode == 1 ? idx[n] : n;
  out[n] = T[s];
  out[(size_t)N + n] = T[(size_t)S + s];
  out[2 * (size_t)N + n] = T[2 * (size_t)S + s];
  if (idx_out) idx_out[n] = s;
}

// Sharded filter: rows of the resampled set this rank produced.  Global slot k in [first, first +
// count) was filled by the local source idx[k - kbase]; a row is the 12 floats of T (3 x float4).
__global__ void __launch_bounds__(PF_BLOCK) pf_shard_export_kernel(const float4* T, const int* idx, int S, long long kbase,
                                                                   long long first, long long count, int use_idx,
                                                                   float4* rows)
{
  const long long i = (long long)blockIdx.x * PF_BLOCK + threadIdx.x;
  if (i >= count) return;
  const long long k = first + i - kbase;
  const int s = use_idx ? idx[k] : (int)k;
  rows[3 * i] = T[s];
  rows[3 * i + 1] = T[(size_t)S + s];
  rows[3 * i + 2] = T[2 * (size_t)S + s];
}

// Sharded filter: rows received from other ranks become local particles [off, off + count) of the
// set read by the next frame.
__global__ void __launch_bounds__(PF_BLOCK) pf_shard_import_kernel(const float4* rows, int S, int off, int count, float4* dst)
{
  const int i = blockIdx.x * PF_BLOCK + threadIdx.x;
  if (i >= count) return;
  const int n = off + i;
  dst[n] = rows[3 * (size_t)i];
  dst[(size_t)S + n] = rows[3 * (size_t)i + 1];
  dst[2 * (size_t)S + n] = rows[3 * (size_t)i + 2];
}

// Sharded filter: output slots that stay on this rank go straight from T (through idx) to the
// next frame's particle set, plane to plane: slot first + i -> local particle off + i.
__global__ void __launch_bounds__(PF_BLOCK) pf_shard_keep_kernel(const float4* T, const int* idx, int S, long long kbase,
                                                                 long long first, int count, int off, float4* dst)
{
  const int i = blockIdx.x * PF_BLOCK + threadIdx.x;
  if (i >= count) return;
  const int s = idx[first + i - kbase];
  const int n = off + i;
  dst[n] = T[s];
  dst[(size_t)S + n] = T[(size_t)S + s];
  dst[2 * (size_t)S + n] = T[2 * (size_t)S + s];
}

// Sharded filter, peer-memory rebalance.  Every output slot this rank produced (slot k = kbase + i, filled by the local
// source idx[i]) belongs to rank k / N as its local slot k % N.  Nothing is copied for a slot that stays on its rank: the
// propagated set T is the next frame's source set, and the slot's entry of lead[] simply names the source row.  A slot of
// ANOTHER rank gets its row through NVLink peer stores into the receive area behind that rank's rows (plane entries
// [N, 2N)), and lead[] points there.  Resampling replicates, so a run of slots with the same source inside one
// destination's range travels as ONE row (stored at the run's first slot) and every slot of the run points to it; a slot
// that continues a run has a source with >= 2 offspring, so the run starts at kend[s - 1] (the rank's first slot for
// s = 0), clamped to the destination's first slot.  In a balanced frame the pass therefore moves 8 bytes per slot (idx
// in, lead out) instead of 100 (idx, row in, row out).  bases[d] = rank d's allocation [pose0 planes | pose1 planes |
// lead], tsel = which pose buffer holds T (the same on every rank), S = plane stride = 2 N.
__global__ void __launch_bounds__(PF_BLOCK) pf_shard_scatter_kernel(const float4* T, const int* idx, const int* kend, int N, int S,
                                                                    long long kbase, long long count, float4* const* bases,
                                                                    int tsel, int me)
{
  const long long stride = (long long)gridDim.x * PF_BLOCK;
  for (long long i = (long long)blockIdx.x * PF_BLOCK + threadIdx.x; i < count; i += stride) {
    const long long k = kbase + i;
    const int d = (int)(k / N);
    const long long d0 = (long long)d * N;
    const int n = (int)(k - d0);
    const int s = idx[i];
    float4* base = bases[d];
    int* lead = reinterpret_cast<int*>(base + 6 * (size_t)S);
    if (d == me) {
      lead[n] = s;
    } else {
      float4* dst = base + (tsel ? 3 * (size_t)S : 0) + N;            // receive area of the destination's T buffer
      const bool starts = i == 0 || k == d0 || idx[i - 1] != s;
      if (starts) {
        const float4 r0 = T[s], r1 = T[(size_t)S + s], r2 = T[2 * (size_t)S + s];
        dst[n] = r0;
        dst[(size_t)S + n] = r1;
        dst[2 * (size_t)S + n] = r2;
        lead[n] = N + n;
      } else {
        const long long fs = s > 0 ? (long long)kend[s - 1] : kbase;
        lead[n] = N + (int)((fs > d0 ? fs : d0) - d0);
      }
    }
  }
}

// The same rebalance with the cross-rank rows PACKED: with run de-duplication the sparse form above stores a row only at
// the first slot of a run, i.e. isolated 16-byte NVLink writes (measured on 8 B200: 1.5-1.7 ms per frame for the rank that
// holds most of the weight, a third of the link rate).  Here a CTA takes PF_BLOCK consecutive global slots -- shards are
// a multiple of PF_BLOCK, so the chunk has one destination -- numbers the runs that start in it (a chunk boundary also
// starts one: at most one extra row per PF_BLOCK slots), reserves that many consecutive places in the destination's
// receive area with one atomic on a per-destination cursor, collects the runs' source indices in shared memory and then
// moves the rows with ALL threads: consecutive threads, consecutive remote addresses, three 16-byte planes.  Source rank
// r owns the places [first slot of r in d, ...) of d's receive area: the streams of different ranks cannot collide, and a
// stream is never longer than the number of slots it serves.  Which place a row gets depends on the order in which the
// CTAs reach the atomic; lead[] records it, so the resampled set the next score pass gathers is the same whatever the
// order.  cursor: [world] zeroed before the launch.
__global__ void __launch_bounds__(PF_BLOCK) pf_shard_scatter_dense_kernel(const float4* T, const int* idx, int N, int S, long long kbase,
                                                                          long long count, float4* const* bases, int tsel, int me,
                                                                          unsigned* cursor)
{
  __shared__ int list[PF_BLOCK];
  __shared__ int wsum[PF_BLOCK / 32];
  __shared__ int s_base;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const long long k_first = kbase / PF_BLOCK * PF_BLOCK;
  const long long nchunks = (kbase + count - k_first + PF_BLOCK - 1) / PF_BLOCK;
  for (long long c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const long long kc = k_first + c * PF_BLOCK, k = kc + t;
    const bool valid = k >= kbase && k < kbase + count;
    const int d = (int)(kc / N);                                   // uniform: N is a multiple of PF_BLOCK
    const long long d0 = (long long)d * N, i = k - kbase;
    const int n = (int)(k - d0);
    const int s = valid ? idx[i] : -1;
    float4* base = bases[d];
    int* lead = reinterpret_cast<int*>(base + 6 * (size_t)S);
    if (d == me) {                                                 // stays on this rank: the source row is not moved
      if (valid) lead[n] = s;
      continue;
    }
    const int sp = (valid && i > 0) ? idx[i - 1] : -2;
    const bool starts = valid && (t == 0 || sp != s);
    const unsigned m = __ballot_sync(PF_FULL, starts);
    const int incl_w = __popc(m & (0xFFFFFFFFu >> (31 - lane)));
    if (lane == 0) wsum[warp] = __popc(m);
    __syncthreads();
    int woff = 0, total = 0;
#pragma unroll
    for (int w = 0; w < PF_BLOCK / 32; ++w) { const int v = wsum[w]; woff += w < warp ? v : 0; total += v; }
    const int incl = woff + incl_w;                                // runs started at or before this slot (>= 1 for a valid slot)
    if (t == 0) s_base = (int)atomicAdd(&cursor[d], (unsigned)total);
    if (starts) list[incl - 1] = s;
    __syncthreads();
    const int pos0 = (int)((kbase > d0 ? kbase : d0) - d0) + s_base;   // this rank's stream in d's receive area + the reservation
    if (valid) lead[n] = N + pos0 + incl - 1;
    if (t < total) {
      const int src = list[t];
      float4* dst = base + (tsel ? 3 * (size_t)S : 0) + N + pos0 + t;
      const float4 r0 = T[src], r1 = T[(size_t)S + src], r2 = T[2 * (size_t)S + src];
      dst[0] = r0;
      dst[(size_t)S] = r1;
      dst[2 * (size_t)S] = r2;
    }
    __syncthreads();                                               // list / wsum / s_base are reused by the next chunk
  }
}

// Sharded filter: wait until the message of exchange `seq` has arrived from every rank (or from
// `only` alone), sent with peer stores by pf_shard_publish, then -- if host_out is given -- hand the
// table to the host through pinned memory: [world][PF_XMAIL_WORDS] words followed by the sequence
// number (or -1 after ~2 s without a complete set: a peer is gone).  Without host_out it is a
// stream-ordered barrier across the ranks.
__global__ void pf_shard_wait_kernel(const long long* mail, int slot, int world, int only, long long seq, long long* host_out)
{
  const int d = threadIdx.x;
  bool ok = true;
  if (d < world && (only < 0 || d == only)) {
    const volatile long long* m = mail + ((size_t)slot * world + d) * PF_XMAIL_WORDS;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (m[PF_XMAIL_WORDS - 1] != seq) {
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 2000000000ull) { ok = false; break; }      // 2 s of wall time (nanosecond timer: clock independent)
    }
    __threadfence_system();
    if (host_out)
      for (int i = 0; i < PF_XMAIL_WORDS; ++i) host_out[d * PF_XMAIL_WORDS + i] = m[i];
  }
  ok = __all_sync(0xffffffffu, ok);
  if (host_out) {
    __threadfence_system();
    __syncwarp();
    if (d == 0) *reinterpret_cast<volatile long long*>(host_out + (size_t)world * PF_XMAIL_WORDS) = ok ? seq : -1ll;
  }
}

// Position-keyed checksums of one tracker's state: [0] the propagated set (PoseParticle, kept iteration), [1] its raw
// weights, [2] the resampled set (newPoseEstimation = rows gathered through idx / lead).  Every 32-bit word is mixed
// with its GLOBAL position (particle or slot number x word number) and the 64-bit results are summed: the sum does not
// depend on how the particles are split over CTAs, handles or GPUs, so the sums of the shards of a sharded filter add
// up to the checksum of the unsharded one exactly when every word sits at the same global position.
__device__ __forceinline__ unsigned long long pf_mix64(unsigned long long pos, unsigned v)
{
  unsigned long long z = pos * 0x9E3779B97F4A7C15ull + ((unsigned long long)v + 1ull) * 0xBF58476D1CE4E5B9ull;
  z ^= z >> 31; z *= 0x94D049BB133111EBull; z ^= z >> 29;
  return z;
}
__global__ void __launch_bounds__(PF_BLOCK) pf_checksum_kernel(const float4* T, const float* w, const float4* src, const int* idx,
                                                               int src_mode, int N, int S, unsigned long long goff, unsigned long long* out)
{
  unsigned long long a = 0, b = 0, c = 0;
  for (long long n = (long long)blockIdx.x * PF_BLOCK + threadIdx.x; n < N; n += (long long)gridDim.x * PF_BLOCK) {
    const unsigned long long g = goff + (unsigned long long)n;
    const int s = src_mode == 1 ? idx[n] : (int)n;
#pragma unroll
    for (int pl = 0; pl < 3; ++pl) {
      const float4 t = T[(size_t)pl * S + n], r = src[(size_t)pl * S + s];
      a += pf_mix64(g * 16 + 4 * pl + 0, __float_as_uint(t.x)) + pf_mix64(g * 16 + 4 * pl + 1, __float_as_uint(t.y)) +
           pf_mix64(g * 16 + 4 * pl + 2, __float_as_uint(t.z)) + pf_mix64(g * 16 + 4 * pl + 3, __float_as_uint(t.w));
      c += pf_mix64(g * 16 + 4 * pl + 0, __float_as_uint(r.x)) + pf_mix64(g * 16 + 4 * pl + 1, __float_as_uint(r.y)) +
           pf_mix64(g * 16 + 4 * pl + 2, __float_as_uint(r.z)) + pf_mix64(g * 16 + 4 * pl + 3, __float_as_uint(r.w));
    }
    b += pf_mix64(g, __float_as_uint(w[n]));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(PF_FULL, a, o); b += __shfl_xor_sync(PF_FULL, b, o); c += __shfl_xor_sync(PF_FULL, c, o);
  }
  if ((threadIdx.x & 31) == 0) { atomicAdd(out, a); atomicAdd(out + 1, b); atomicAdd(out + 2, c); }
}

// Sharded filter: an empty message to every rank (the sending half of a stream-ordered barrier)
__global__ void pf_shard_signal_kernel(long long* const* peers, int world, int rank, int slot, long long seq)
{
  if ((int)threadIdx.x < world) {
    __threadfence_system();
    long long* dst = peers[threadIdx.x] + ((size_t)slot * world + rank) * PF_XMAIL_WORDS;
    *reinterpret_cast<volatile long long*>(dst + PF_XMAIL_WORDS - 1) = seq;
  }
}

#endif  // __CUDACC__
