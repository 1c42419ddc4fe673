// pf_types.h -- device-resident records shared by the kernels (pf_kernels.cuh) and the
// host side of the C ABI (pfmpe_api.cu).  Plain structs (float4 is the only CUDA type).
#pragma once
#include <stdint.h>
#include "pf_tail.cuh"

#define PF_BLOCK 256          /* threads per CTA, batch mode (one CTA per tracker, 3 CTAs / SM)     */
#define PF_GRID_BLOCK 768     /* threads per CTA, grid mode (one CTA per SM: fewest barrier parties) */
#define PF_MAX_BLOCK 768
#define PF_MAX_WARPS (PF_MAX_BLOCK / 32)
#define PF_MAX_GRID 2048      /* capacity of the per-CTA aggregate array (grid mode)  */
#define PF_MAX_ITER 80        /* do/while bound, pose_estimator.cpp:616              */

// Per-tracker persistent state (device memory).
struct pf_tracker_state {
  // FP64 tracker state, used by the device tail (track mode)
  double cur[12], prev[12], pred[12];        // current_pose_, previous_pose_, predicted_pose_ (affine row-major)
  double t_cur, t_prev, t_pred;
  double cov[36];
  uint32_t isi;
  int32_t failed;                            // track mode: set when a frame ends with Flag_Fail 4
  // particle-set bookkeeping
  int32_t src_buf;                           // pose buffer that holds the source set of the next frame
  int32_t src_mode;                          // 0: E[n] = pose[src_buf][n];  1: E[n] = pose[src_buf][idx(n)] via kend (resampled)
  uint32_t reserved0;
  int32_t frames_done;
  int32_t t_buf;                             // pose buffer that holds the last propagated set (PoseParticle)
  int32_t pad;
};

// Result record of one frame (device -> host).
struct pf_frame_result {
  int32_t status;                 // 0 not run, 1 accepted (Flag_Fail 1), 4 rejected (Flag_Fail 4)
  int32_t iterations;             // do/while iterations executed
  int32_t best_iteration;         // iteration whose particles were kept (-1: none scored > 0)
  int32_t resampled;
  float   highest_prob;
  int32_t most_likely_particle_idx;   // first argmax weight (kept iteration)
  int32_t most_likely_pose_idx;       // first argmax resample count
  uint32_t max_count;
  long long sum_fx;               // weight sum, Q.30
  float   winner_pose[12];        // predicted_pose_ leaving the PF block (affine row-major, FP32 particle)
  uint32_t pairs[2 * PF_MAX_LED]; // winner correspondences (LED+1, blob+1), greedy order
  int32_t npairs;
  int32_t gn_iters;               // track mode: PubData.numIter
  float   flag_fail;              // track mode: 1, 1.5 or 4
  int32_t pad;
  double  pose[12];               // track mode: refined pose (current_pose_ after updatePose)
  double  cov[36];                // track mode: pose_covariance_
  double  mean_pose[12];          // optional weighted mean (north_star extra)
  double  mean_cov[36];
};

// Cross-CTA accumulators for one reduction round (grid mode); zero-initialised.
struct pf_accum {
  unsigned long long maxkey;      // (ord(w) << 32) | ~idx          -> atomicMax: max w, first index
  unsigned long long minkey_inv;  // ~((ord(w) << 32) | idx)        -> atomicMax: min w, first index
  long long sum;                  // sum of Q.30 weights
  unsigned long long cntkey;      // (count << 32) | ~idx           -> atomicMax: max count, first index
};

// Per-CTA aggregate of the sign-agnostic prefix statistics (grid mode), written before the
// grid-wide barrier that follows every score pass; 32-byte records.
struct pf_cta_agg {
  long long s, mx, mn, pad;
};

#define PF_INLINE_BLOBS 32
#define PF_XMAIL_RING 4
#define PF_XMAIL_WORDS 32   /* message: up to 31 payload words + the sequence number in the last word */

// grid mode: heavy steps of the current resampling, filled by the whole grid (pf_grid_fill).
// ent = (first source of the step, first slot, number of slots, 0)
struct pf_gfill {
  int n, ktot, kbase, pad;
#if defined(__CUDACC__)
  int4 ent[1];
#else
  int ent[4];
#endif
};

struct pf_kparams {
  int N;                          // particles per tracker
  int n_trackers;
  int tracker_base;               // first tracker handled by blockIdx.y == 0
  int D_max;                      // capacity of the per-frame blob arrays
  int mode;                       // 0: per-frame (host tail), 1: device-resident tracking (device tail),
                                  // 2: replay of a tracked sequence: PF updates only, constants from fc_log
  int n_frames;                   // frames processed by this launch
  int frame0;                     // index of the first frame in the uploaded sequence (track mode)
  int total_frames;               // frames uploaded (track mode)
  int coop;                       // 1: grid-wide cooperation (gridDim.x > 1)
  int want_mean;                  // compute weighted mean / covariance
  int cta_chunk;                  // particles owned by one CTA (multiple of 32 * warps per CTA); warp w owns the w-th part
  int pstride;                    // plane stride of the pose buffers in float4: N, or 2 N for a sharded filter (receive area)
  // particle state
  float4* pose0; float4* pose1;   // [tracker][3 planes][pstride]
  float*  w;                      // [tracker][N] raw scores of the kept iteration
  int*    kend;                   // [tracker][N] cumulative offspring count K(M_i)
  int*    idx;                    // [tracker][N] Particle_index of output slot n (:678-681), written by the resampling stage
  pf_cta_agg* cta_agg;            // [PF_MAX_GRID] (grid mode)
  pf_accum* accum;                // [3] rotating slots (grid mode); [3]: maxkey = next round mod 3, minkey_inv = barriers done, cntkey = arrival counter
  // per-tracker records
  pf_frame_const* fconst;         // [tracker]
  pf_frame_const* fc_log;         // [tracker][total_frames] constants of every tracked frame (optional)
  pf_tracker_state* state;        // [tracker]
  pf_frame_result* result;        // per-frame mode: [tracker]; track mode: [tracker][total_frames]
  // inputs
  const double* blobs;            // per-frame mode: [tracker][D_max][2]; track mode: [tracker][total_frames][D_max][2]
  const int* blob_count;          // track mode: [tracker][total_frames]
  const double* times;            // track mode: [total_frames] time_to_predict
  const uint32_t* seeds;          // track mode: [tracker][total_frames]
  // constants
  const float* obj_f;             // [tracker][PF_MAX_LED*3]
  const double* obj_d;            // [tracker][PF_MAX_LED*3]
  float K9f[9];
  int kstd;                       // K has the usual zero pattern (pf_k_is_std)
  double K9d[9];
  pf_score_cfg score;
  double min_ang, max_ang, min_tr, max_tr;   // minAngularNoise.. (track mode recomputes the ranges per frame)
  const uint32_t* rng_tab;        // [4*256]
  unsigned long long* timeline;   // optional [16] globaltimer stamps of the last frame (CTA 0, thread 0)
  // per-frame API, one tracker, few detections: the frame's inputs travel in the parameter block
  // (no host->device copy before the launch)
  int inline_in;
  int done_seq;                   // != 0: the result record lives in pinned host memory; published last, after a system fence
  int gn_cap;                     // track mode: > 0 caps the device Gauss-Newton (gn_iters = -2: the host finishes the object)
  int split_tail;                 // track mode, one frame per launch, batch shape: the FP64 tail runs as pf_track_tail_kernel (one warp per filter)
  int prefetch;                   // score pass: prefetch the next step's source rows into L1
  int tail_count;                 // ... over the filters [tracker_base, tracker_base + tail_count)
  pf_frame_const fc_inline;
  double blobs_inline[2 * PF_INLINE_BLOBS];
  float dist[5];                  // plumb_bob (k1, k2, p1, p2, k3), used by the V = 2 kernel (PFMPE_FLAG_DISTORT)
  pf_gfill* gfill;                // grid mode: list of the heavy steps of the current resampling
  int gfill_cap;                  // entries it can hold
  // sharded single filter (modes 3-5): this handle owns global particles [goff, goff + N) of N_global
  int goff, N_global;
  int pass_it;                    // mode 3: iteration to run
  int iters_total;                // modes 4/5: do/while iterations executed (all shards)
  int q_local;                    // mode 5: local index of the winner
  int idx_cap;                    // capacity of idx[] (slots this shard's sources may own)
  int lead_mode;                  // mode 4: the rebalance leaves one row per run of equal sources + lead[] (pf_shard_scatter_kernel)
  const int* lead;                // sharded: [N] local slot that holds the row of slot n (gather indirection of the score pass)
  long long S_fx;                 // modes 4/5: global weight sum, Q.30
  long long rank_excl[3], rank_total[3];   // mode 4: (sum, max, min) aggregate of the lower ranks / of all ranks
  long long* shard_partial;       // [8] per-stage results of this shard (device)
  pf_cta_agg* warp_agg_g;         // [grid][PF_MAX_WARPS] per-warp prefix aggregates of the last score pass
  // peer mailboxes (optional): the stage kernel itself delivers its partial to every rank, NVLink peer stores
  long long* const* xmail_peers;  // [xworld] base of every rank's mailbox: [PF_XMAIL_RING][xworld][PF_XMAIL_WORDS] int64
  int xworld, xrank, xslot;
  long long xseq;                 // sequence number of this exchange (word 7 of the message, written last)
};
