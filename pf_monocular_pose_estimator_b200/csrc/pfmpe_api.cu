// pfmpe_api.cu -- host side of the C ABI declared in include/pfmpe.h.
// Owns device memory, the stream and the launch geometry; all arithmetic of the hot path is in
// the sm_100a kernels of pf_kernels.cuh.  There is no CPU execution path: every entry point that
// needs the device fails with PFMPE_ERR_NO_DEVICE / PFMPE_ERR_CUDA if it is not usable.
#include <cuda_runtime.h>

#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "../../include/pfmpe.h"
#include "pf_kernels.cuh"
#include "pf_meancov.cuh"
#include "pf_roi.cuh"
#include "pf_p3p.cuh"

namespace {

struct TrackSeq {
  double* blobs = nullptr;
  int* counts = nullptr;
  double* times = nullptr;
  uint32_t* seeds = nullptr;
  pf_frame_result* result = nullptr;
  pf_frame_const* fc_log = nullptr;   // constants of every tracked frame (replay)
  int n_frames = 0;
  bool owned = false;
  int next_frame = -1;     // frame whose constants are currently staged in fconst (track mode)
  int last_frame = -1;     // last frame processed by track_run
  int tracked_upto = -1;   // highest frame whose constants are in fc_log
};

// Persistent host workers for the per-tracker FP64 tail of a batch (predictPose before, Gauss-Newton after the PF update)
// and the staging / unpacking of its records: independent per tracked object, so a batch of thousands is spread over the
// host cores (the reference's `for objectNumber` loop, :89, is serial).  Workers sleep on a condition variable between
// calls (no spinning: several ranks may share one host); the caller takes a share of the work itself.
class HostPool {
 public:
  explicit HostPool(int workers)
  {
    for (int i = 0; i < workers; ++i) th_.emplace_back([this, i]() { loop(i); });
  }
  ~HostPool()
  {
    { std::lock_guard<std::mutex> g(m_); stop_ = true; ++gen_; }
    cv_.notify_all();
    for (auto& t : th_) t.join();
  }
  int size() const { return (int)th_.size(); }
  // body(i) for i in [0, n): contiguous chunks, one per worker + one for the caller
  void run(int n, const std::function<void(int)>& body)
  {
    const int parts = (int)th_.size() + 1, per = (n + parts - 1) / parts;
    {
      std::lock_guard<std::mutex> g(m_);
      body_ = &body; n_ = n; per_ = per; pending_ = (int)th_.size(); ++gen_;
    }
    cv_.notify_all();
    for (int i = (parts - 1) * per; i < n; ++i) body(i);        // the caller's chunk (the last one)
    std::unique_lock<std::mutex> g(m_);
    done_.wait(g, [this]() { return pending_ == 0; });
    body_ = nullptr;
  }
  // the same for a few items of very uneven cost (a Gauss-Newton of 70 or of 500 iterations): every thread takes the next
  // item as it becomes free
  void run_uneven(int n, const std::function<void(int)>& body)
  {
    {
      std::lock_guard<std::mutex> g(m_);
      next_.store(0);
      body_ = &body; n_ = n; per_ = -1; pending_ = (int)th_.size(); ++gen_;
    }
    cv_.notify_all();
    for (int i; (i = next_.fetch_add(1)) < n;) body(i);
    std::unique_lock<std::mutex> g(m_);
    done_.wait(g, [this]() { return pending_ == 0; });
    body_ = nullptr;
  }

 private:
  void loop(int id)
  {
    unsigned long long seen = 0;
    for (;;) {
      const std::function<void(int)>* body;
      int n, per;
      {
        std::unique_lock<std::mutex> g(m_);
        cv_.wait(g, [&]() { return gen_ != seen; });
        seen = gen_;
        if (stop_) return;
        body = body_; n = n_; per = per_;
      }
      if (per < 0) for (int i; (i = next_.fetch_add(1)) < n;) (*body)(i);
      else for (int i = id * per; i < n && i < (id + 1) * per; ++i) (*body)(i);
      {
        std::lock_guard<std::mutex> g(m_);
        if (--pending_ == 0) done_.notify_one();
      }
    }
  }
  std::vector<std::thread> th_;
  std::mutex m_;
  std::condition_variable cv_, done_;
  const std::function<void(int)>* body_ = nullptr;
  int n_ = 0, per_ = 0, pending_ = 0;
  std::atomic<int> next_{0};
  unsigned long long gen_ = 0;
  bool stop_ = false;
};

// host-side tracked object (previous_pose_, current_pose_, their times, it_since_initialized_, covariance)
struct HostTrack {
  pf_aff prev, cur;
  double t_prev = 0, t_cur = 0;
  uint32_t isi = 0;
  double cov[36] = {0};
  bool valid = false;
};

}  // namespace

struct pfmpe_handle {
  pfmpe_config cfg{};
  pfmpe_pf_params prm{};
  bool have_markers = false, have_camera = false, have_params = false;
  std::vector<char> particles_set;
  std::vector<double> markers;   // [tracker][PF_MAX_LED*3]
  double K[9]{};
  double D5[5]{};
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool timing_valid = false;
  int done_seq = 0;
  uint64_t launches = 0;
  int num_sms = 0;
  int force_mode = 0;            // 0 auto, 1 batch, 2 grid (cfg.reserved[0])
  int grid_cap = 0;              // cfg.reserved[1]
  int info[4] = {0, 0, PF_BLOCK, 0};
  int batch_nt = -1, batch_block = PF_BLOCK;   // cached block-size choice of the batch shape
  const void* batch_kernel = nullptr;
  // device
  float4 *pose0 = nullptr, *pose1 = nullptr, *gather_tmp = nullptr;
  float* w = nullptr;
  int *kend = nullptr, *idx = nullptr, *gather_idx = nullptr;
  pf_cta_agg* cta_agg = nullptr;
  pf_accum* accum = nullptr;
  pf_gfill* gfill = nullptr;
  unsigned* roi_keys = nullptr;   // K8: [tracker][4] ordered-float keys, then the floats
  float* roi_out = nullptr;
  float* h_roi = nullptr;
  int* lead = nullptr;                  // sharded: [N] gather indirection, behind the pose planes (one IPC allocation)
  size_t pstride = 0;                   // plane stride of pose0 / pose1 in float4: N, or 2 N for a sharded filter
  bool shard_alloc = false;             // pose0 / pose1 / lead are one allocation (freed through pose0)
  int sh_src = 0;                       // sharded: host mirror of the source-buffer selector (stage calls are issued in order)
  unsigned char* p3p_buf = nullptr;     // K9 scratch: detections, sorted copy, bearing vectors, counters, histogram
  size_t p3p_bytes = 0;
  double2* p3p_roots = nullptr;         // K9: roots of the quartics of one slab of problems (two-kernel form of large votes)
  size_t p3p_roots_bytes = 0;
  double* p3p_io = nullptr;             // batched solver scratch
  size_t p3p_io_bytes = 0;
  double *mc_part = nullptr, *mc_mean = nullptr;   // K7 (PFMPE_FLAG_MEAN_COV): per-block partials, mean poses
  int mc_blocks = 0;
  int gfill_cap = 0;
  unsigned char* inbox = nullptr;       // [fconst x n_trackers | blobs x n_trackers x max_blobs x 2]
  pf_frame_const* fconst = nullptr;     // = inbox
  double* blobs = nullptr;              // = inbox + ...
  size_t inbox_bytes = 0, blobs_off = 0;
  pf_tracker_state* state = nullptr;
  pf_frame_result* result = nullptr;
  float* obj_f = nullptr;
  double* obj_d = nullptr;
  uint32_t* rng_tab = nullptr;
  unsigned long long* timeline = nullptr;
  bool timeline_on = false;
  // pinned host mirrors
  unsigned char* h_inbox = nullptr;
  pf_frame_result* h_result = nullptr;
  pf_tracker_state* h_state = nullptr;
  std::vector<long long> last_sum_fx;
  bool last_was_track = false;
  TrackSeq seq;
  std::vector<HostTrack> host_trk;   // host mirror of the tracked objects (pfmpe_track_frame_host)
  int gn_cap = 0;                    // launch_frames: Gauss-Newton cap of the next track-mode launch (live frames only)
  TrackSeq live;                     // one-frame sequence of pfmpe_track_frame_device_batch (device buffers, created on first use)
  unsigned char* h_live = nullptr;   // pinned staging of its inputs: blobs [T][D][2] | time | counts [T] | seeds [T]
  pf_frame_result* h_live_res = nullptr;   // pinned: the T result records of the frame
  int* h_late_hdr = nullptr;               // pinned, written by pf_track_late_kernel: [0] count, [4..] tracker list
  pf_frame_result* h_late_rec = nullptr;   // pinned, written by pf_track_late_kernel: the records of those trackers
  cudaEvent_t ev_late = nullptr;
  // sharded single filter (pfmpe_shard_*): this handle owns global particles [goff, goff + N) of n_global
  bool sharded = false;
  int goff = 0, n_global = 0;
  pf_cta_agg* warp_agg_g = nullptr;
  long long* shard_partial = nullptr;   // default target of the per-stage partial results
  long long* xmail = nullptr;           // this rank's mailbox [PF_XMAIL_RING][world][PF_XMAIL_WORDS]
  long long** xmail_peers_dev = nullptr;
  long long* h_xtable = nullptr;        // pinned: [world][PF_XMAIL_WORDS] + sequence word
  int xworld = 0, xrank = 0;
  long long xseq = 0;
  float4** peers_dev = nullptr;         // [peer_world] next-frame particle planes of every rank (peer memory)
  unsigned* scatter_cursor = nullptr;   // [peer_world] places reserved per destination by the packed rebalance
  int peer_world = 0;
  std::vector<void*> ipc_opened;
  std::unique_ptr<HostPool> pool;       // host workers of the batched host tail (created on first use)
  uint64_t shard_passes = 0;            // score passes launched by pfmpe_shard_frame
  unsigned long long* cksum = nullptr;  // pfmpe_state_checksum scratch
  std::vector<pfmpe_frame_in> in_scratch;
  std::vector<pfmpe_frame_out> out_scratch;
  std::string err;
};

static void host_parallel_for(pfmpe_handle* h, int n, const std::function<void(int)>& body);

namespace {

thread_local std::string g_create_err;

#define PF_CUDA(h, call)                                                                         \
  do {                                                                                           \
    cudaError_t e_ = (call);                                                                     \
    if (e_ != cudaSuccess) {                                                                     \
      (h)->err = std::string(#call) + ": " + cudaGetErrorString(e_);                             \
      return PFMPE_ERR_CUDA;                                                                     \
    }                                                                                            \
  } while (0)

int fail(pfmpe_handle* h, int code, const char* msg)
{
  if (h) h->err = msg;
  return code;
}

template <typename T>
cudaError_t dev_alloc(T** p, size_t n)
{
  return cudaMalloc((void**)p, n * sizeof(T));
}

void free_all(pfmpe_handle* h)
{
  cudaFree(h->pose0); if (!h->shard_alloc) cudaFree(h->pose1); cudaFree(h->gather_tmp); cudaFree(h->w); cudaFree(h->kend);
  cudaFree(h->idx); cudaFree(h->gather_idx); cudaFree(h->cta_agg); cudaFree(h->accum);
  cudaFree(h->inbox); cudaFree(h->state); cudaFree(h->result); cudaFree(h->obj_f); cudaFree(h->obj_d);
  for (void* q : h->ipc_opened) cudaIpcCloseMemHandle(q);
  cudaFree(h->xmail); cudaFree(h->xmail_peers_dev);
  if (h->h_xtable) cudaFreeHost(h->h_xtable);
  cudaFree(h->peers_dev); cudaFree(h->scatter_cursor); cudaFree(h->mc_part); cudaFree(h->mc_mean); cudaFree(h->roi_keys); cudaFree(h->roi_out);
  if (h->h_roi) cudaFreeHost(h->h_roi);
  cudaFree(h->p3p_buf); cudaFree(h->p3p_roots); cudaFree(h->p3p_io); cudaFree(h->cksum);
  cudaFree(h->rng_tab); cudaFree(h->timeline); cudaFree(h->warp_agg_g); cudaFree(h->shard_partial); cudaFree(h->gfill);
  if (h->seq.owned) { cudaFree(h->seq.blobs); cudaFree(h->seq.counts); cudaFree(h->seq.times); cudaFree(h->seq.seeds); }
  cudaFree(h->seq.result); cudaFree(h->seq.fc_log);
  cudaFree(h->live.blobs); cudaFree(h->live.counts); cudaFree(h->live.times); cudaFree(h->live.seeds); cudaFree(h->live.result);
  if (h->h_live) cudaFreeHost(h->h_live);
  if (h->h_live_res) cudaFreeHost(h->h_live_res);
  if (h->h_late_hdr) cudaFreeHost(h->h_late_hdr);
  if (h->h_late_rec) cudaFreeHost(h->h_late_rec);
  if (h->ev_late) cudaEventDestroy(h->ev_late);
  if (h->h_inbox) cudaFreeHost(h->h_inbox);
  if (h->h_result) cudaFreeHost(h->h_result);
  if (h->h_state) cudaFreeHost(h->h_state);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
}

void fill_kparams(const pfmpe_handle* h, pf_kparams& p)
{
  std::memset(&p, 0, sizeof p);
  p.N = h->cfg.n_particles;
  p.n_trackers = h->cfg.n_trackers;
  p.D_max = h->cfg.max_blobs;
  p.want_mean = (h->cfg.flags & PFMPE_FLAG_MEAN_COV) ? 1 : 0;
  p.pstride = (int)h->pstride;
  p.pose0 = h->pose0; p.pose1 = h->pose1; p.w = h->w; p.kend = h->kend; p.idx = h->idx;
  p.cta_agg = h->cta_agg; p.accum = h->accum; p.gfill = h->gfill; p.gfill_cap = h->gfill_cap;
  p.fconst = h->fconst; p.state = h->state;
  p.obj_f = h->obj_f; p.obj_d = h->obj_d;
  for (int i = 0; i < 9; ++i) { p.K9d[i] = h->K[i]; p.K9f[i] = (float)h->K[i]; }
  p.kstd = pf_k_is_std(p.K9f) ? 1 : 0;
  for (int i = 0; i < 5; ++i) p.dist[i] = (float)h->D5[i];
  p.score.tol = (float)h->prm.back_projection_pixel_tolerance;
  p.score.inv_tol = (float)(1.0 / h->prm.back_projection_pixel_tolerance);
  p.score.tol_pf = (float)h->prm.back_projection_pixel_tolerance_PF;
  unsigned m = 0;
  for (int i = 0; i < 5; ++i) if (h->prm.marker_downgrade[i]) m |= 1u << i;
  p.score.dg_mask = m;
  pf_score_cfg_finish(p.score, h->cfg.n_leds);
  p.min_ang = h->prm.min_angular_noise; p.max_ang = h->prm.max_angular_noise;
  p.min_tr = h->prm.min_transition_noise; p.max_tr = h->prm.max_transition_noise;
  p.rng_tab = h->rng_tab;
  p.goff = h->sharded ? h->goff : 0;
  p.N_global = h->sharded ? h->n_global : h->cfg.n_particles;
  p.idx_cap = p.N_global;
  p.warp_agg_g = h->warp_agg_g;
  p.shard_partial = h->shard_partial;
  p.lead = h->sharded ? h->lead : nullptr;
  p.lead_mode = (h->sharded && h->peers_dev) ? 1 : 0;
  p.timeline = h->timeline_on ? h->timeline : nullptr;
}

static_assert(sizeof(pf_kparams) <= 4096, "kernel parameter block must stay within the classic 4 KB limit");
typedef void (*frames_kernel_t)(const pf_kparams);

template <int SH>
frames_kernel_t kernel_for_sh(int L)
{
  switch (L) {
#ifndef PF_DEV_ONLY_L5   /* development builds (make DEV=1) instantiate the 5-LED kernels only: 8x shorter compile */
    case 1: return pf_frames_kernel<1, SH>;
    case 2: return pf_frames_kernel<2, SH>;
    case 3: return pf_frames_kernel<3, SH>;
    case 4: return pf_frames_kernel<4, SH>;
    case 6: return pf_frames_kernel<6, SH>;
    case 7: return pf_frames_kernel<7, SH>;
    case 8: return pf_frames_kernel<8, SH>;
#endif
    case 5: return pf_frames_kernel<5, SH>;
    default: return nullptr;
  }
}

typedef void (*tail_kernel_t)(const pf_kparams, int);
tail_kernel_t tail_kernel_for(int L)
{
  switch (L) {
#ifndef PF_DEV_ONLY_L5
    case 1: return pf_track_tail_kernel<1>; case 2: return pf_track_tail_kernel<2>; case 3: return pf_track_tail_kernel<3>;
    case 4: return pf_track_tail_kernel<4>; case 6: return pf_track_tail_kernel<6>; case 7: return pf_track_tail_kernel<7>;
    case 8: return pf_track_tail_kernel<8>;
#endif
    case 5: return pf_track_tail_kernel<5>;
    default: return nullptr;
  }
}

frames_kernel_t kernel_for(int L, int variant)
{
  return variant == 1 ? kernel_for_sh<1>(L) : (variant == 2 ? kernel_for_sh<2>(L) : (variant == 3 ? kernel_for_sh<3>(L) : kernel_for_sh<0>(L)));
}

// per-launch arguments of the sharded-filter stages (modes 3-5)
struct ShardArgs {
  int pass_it = 0, iters_total = 0, q_local = 0;
  long long S_fx = 0;
  long long excl[3] = {0, 0, 0}, total[3] = {0, 0, 0};
  long long* partial = nullptr;
  bool publish = false;        // mode 5: send the winner record to every rank's mailbox
};

// K7: weighted mean pose + covariance of trackers [t0, t0+nt) from the kept iteration's particles and
// scores (two streaming passes + two one-block combines), written into the result records.
int launch_meancov(pfmpe_handle* h, int t0, int nt)
{
  pf_kparams p;
  fill_kparams(h, p);
  p.tracker_base = t0;
  p.result = h->result;
  const dim3 grid((unsigned)h->mc_blocks, (unsigned)nt, 1);
  double* mean = h->mc_mean + (size_t)t0 * 12;
  pf_meancov_pass1<<<grid, PF_MC_THREADS, 0, h->stream>>>(p, h->mc_part, 0ll);
  pf_meancov_mean<<<nt, 32, 0, h->stream>>>(p, h->mc_part, h->mc_blocks, mean);
  pf_meancov_pass2<<<grid, PF_MC_THREADS, 0, h->stream>>>(p, mean, h->mc_part, 0ll);
  pf_meancov_cov<<<nt, 32, 0, h->stream>>>(p, h->mc_part, h->mc_blocks);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 4;
  return PFMPE_OK;
}

// per-frame API, latency path: inputs inside the parameter block, result record written to pinned host memory
struct InlineArgs {
  const pf_frame_const* fc = nullptr;
  const double* blobs = nullptr;   // D x 2, D <= PF_INLINE_BLOBS
  int D = 0;
  pf_frame_result* result = nullptr;   // device-accessible (mapped pinned) array indexed by tracker
  int done_seq = 0;                    // != 0: publish the record with this sequence number (host polls)
};

// Launch the frame kernel for trackers [t0, t0+nt).  mode 0: per-frame, 1: track, 2: replay,
// 3/4/5: one stage of the sharded filter.
int launch_frames(pfmpe_handle* h, int mode, int t0, int nt, int frame0, int n_frames, const ShardArgs* sa = nullptr,
                  const InlineArgs* ia = nullptr)
{
  const int N = h->cfg.n_particles;
  const bool distort = (h->cfg.flags & PFMPE_FLAG_DISTORT) != 0;
  if (distort && mode != 0) return fail(h, PFMPE_ERR_UNSUPPORTED, "PFMPE_FLAG_DISTORT: per-frame API only (the Gauss-Newton tail needs undistorted points)");
  frames_kernel_t k = kernel_for(h->cfg.n_leds, mode >= 3 ? 1 : (distort ? 2 : 0));
  if (!k) return fail(h, PFMPE_ERR_UNSUPPORTED, "n_leds must be 1..8");
  const size_t smem = pf_blob_smem_bytes(h->cfg.max_blobs);
  bool grid_mode;
  if (h->force_mode == 1 || h->force_mode == 3) grid_mode = false;
  else if (h->force_mode == 2) grid_mode = true;
  else grid_mode = (nt == 1) ? (N > 1024) : (N > 16384);
  if (!grid_mode && mode < 3 && !distort) k = kernel_for(h->cfg.n_leds, 3);   // batch shape: the 64-register build
  if (smem > 48 * 1024)   // candidate lists of a large max_blobs: opt in to more than the default 48 KB of dynamic shared memory
    PF_CUDA(h, cudaFuncSetAttribute((const void*)k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));

  pf_kparams p;
  fill_kparams(h, p);
  p.mode = mode;
  p.frame0 = frame0;
  p.n_frames = n_frames;
  p.gn_cap = mode == 1 ? h->gn_cap : 0;
  // A tracked frame of a batch of filters, one frame per launch: the FP64 tail as a second kernel, one warp per filter
  // (pf_track_tail_kernel; PFMPE_SPLIT_TAIL=0 keeps it inside the frame kernel).  Sequences of several frames per launch
  // keep the fused form: there the tail's result feeds the next frame inside the same launch.
  static const bool split_off = [] { const char* e = std::getenv("PFMPE_SPLIT_TAIL"); return e && std::atoi(e) == 0; }();
  const bool split_tail = mode == 1 && n_frames == 1 && nt > 1 && !grid_mode && !split_off;
  p.split_tail = split_tail ? 1 : 0;
  p.tail_count = nt;
  static const bool prefetch_off = [] { const char* e = std::getenv("PFMPE_PREFETCH"); return e && std::atoi(e) == 0; }();   // A/B
  p.prefetch = prefetch_off ? 0 : 1;
  if (sa) {
    p.pass_it = sa->pass_it; p.iters_total = sa->iters_total; p.q_local = sa->q_local; p.S_fx = sa->S_fx;
    for (int i = 0; i < 3; ++i) { p.rank_excl[i] = sa->excl[i]; p.rank_total[i] = sa->total[i]; }
    if (sa->partial) p.shard_partial = sa->partial;
    if (h->xmail_peers_dev && (mode == 3 || mode == 4 || (mode == 5 && sa->publish))) {   // deliver the partial to every rank's mailbox
      h->xseq += 1;
      p.xmail_peers = h->xmail_peers_dev; p.xworld = h->xworld; p.xrank = h->xrank;
      p.xseq = h->xseq; p.xslot = (int)(h->xseq % PF_XMAIL_RING);
    }
    if (mode == 5) grid_mode = false;     // the winner record is one warp of work
  }
  if (mode == 1 || mode == 2) {
    p.blobs = h->seq.blobs; p.blob_count = h->seq.counts; p.times = h->seq.times; p.seeds = h->seq.seeds;
    p.result = h->seq.result; p.total_frames = h->seq.n_frames; p.fc_log = h->seq.fc_log;
  } else {
    p.blobs = h->blobs; p.result = h->result; p.total_frames = 1;
  }
  if (ia) {
    if (ia->result) p.result = ia->result;
    p.done_seq = ia->done_seq;
    if (ia->fc) {
      p.inline_in = 1;
      p.fc_inline = *ia->fc;
      std::memcpy(p.blobs_inline, ia->blobs, sizeof(double) * 2 * (size_t)ia->D);
    }
  }
  // CUDA events bracket the launches of the sequence modes (pfmpe_last_launch_ms); the per-frame
  // latency path does without them
  const bool timed = mode == 1 || mode == 2;
  if (timed) PF_CUDA(h, cudaEventRecord(h->ev0, h->stream));
  if (!grid_mode) {
    p.coop = 0;
    p.tracker_base = t0;
    // Block size of the one-CTA-per-filter shape: the CTAs of a batch run in waves over the resident
    // slots (SMs x CTAs per SM), so the size that needs the fewest (waves x steps per thread) wins:
    // e.g. 512 filters on 148 SMs fit one wave of 192-thread CTAs (4 per SM) but need two of 256-thread ones.
    int block = PF_BLOCK;
    if (nt > 1 && mode != 5 && h->batch_nt == nt && h->batch_kernel == (const void*)k) block = h->batch_block;
    else if (nt > 1 && mode != 5) {
      long long best = -1, cost256 = -1;
      for (int b = PF_BLOCK; b >= 128; b -= 32) {
        int per_sm = 0;
        PF_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, b, smem));
        if (per_sm < 1) continue;
        const long long slots = (long long)per_sm * h->num_sms;
        const long long cost = ((nt + slots - 1) / slots) * ((N + b - 1) / b);
        if (b == PF_BLOCK) cost256 = cost;
        if (best < 0 || cost < best) { best = cost; block = b; }
      }
      if (cost256 > 0 && best * 10 > cost256 * 9) block = PF_BLOCK;   // the model is coarse: only a clear win changes the default
      if (const char* e = std::getenv("PFMPE_BATCH_BLOCK")) { const int b = std::atoi(e); if (b >= 32 && b <= PF_BLOCK && b % 32 == 0) block = b; }
      h->batch_nt = nt; h->batch_kernel = (const void*)k; h->batch_block = block;
    }
    // Cluster-pair shape (opt-in: cfg.reserved[0] = 3 or PFMPE_CLUSTER=1): every filter is split over a thread-block
    // cluster of two 128-thread CTAs that reduce through distributed shared memory -- twice as many, half as large
    // units for batches that leave the SMs unevenly loaded (512 filters on 148 SMs).  Measured on one B200
    // (profiles/r02_cluster_ab.txt) it is a wash -- 512 filters 0.1136 -> 0.1121 ms, 4096 filters 0.723 -> 0.758 ms: what
    // a small batch loses is not the 4-versus-3 CTAs per SM but the phase lock of a single wave (every CTA stages its
    // detections, then every CTA reduces, then every CTA writes its record at the same time) -- so it is not the default.
    bool pair = h->force_mode == 3 && mode <= 2 && !distort;
    if (const char* e = std::getenv("PFMPE_CLUSTER")) pair = std::atoi(e) != 0 && nt >= 1 && mode <= 2 && !distort && h->force_mode != 1;
    if (pair) {
      const int cblock = 128;
      p.coop = 2;
      p.cta_chunk = ((N + 1) / 2 + cblock - 1) / cblock * cblock;
      h->info[0] = 2; h->info[1] = nt; h->info[2] = cblock; h->info[3] = 2;
      cudaLaunchConfig_t lc;
      std::memset(&lc, 0, sizeof lc);
      lc.gridDim = dim3(2, (unsigned)nt, 1);
      lc.blockDim = dim3(cblock, 1, 1);
      lc.dynamicSmemBytes = smem;
      lc.stream = h->stream;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      lc.attrs = at;
      lc.numAttrs = 1;
      void* args[] = {(void*)&p};
      PF_CUDA(h, cudaLaunchKernelExC(&lc, (const void*)k, args));
      h->launches += 1;
    } else {
    p.cta_chunk = (N + block - 1) / block * block;
    dim3 grid(1, (unsigned)nt, 1);
    h->info[0] = 1; h->info[1] = nt; h->info[2] = block; h->info[3] = 0;
    k<<<grid, block, smem, h->stream>>>(p);
    PF_CUDA(h, cudaGetLastError());
    h->launches += 1;
    }
  } else {
    int block = PF_GRID_BLOCK;
    if (const char* e = std::getenv("PFMPE_GRID_BLOCK")) {   // experiment: the 64-register build, 256-thread CTAs, 4 per SM
      if (std::atoi(e) == 256 && mode < 3 && !distort) { block = 256; k = kernel_for(h->cfg.n_leds, 3); }
    }
    const int warps = block / 32;
    int per_sm = 0;
    PF_CUDA(h, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k, block, smem));
    int cap = per_sm * h->num_sms;
    if (cap < 1) return fail(h, PFMPE_ERR_CUDA, "kernel cannot be made resident");
    if (cap > PF_MAX_GRID) cap = PF_MAX_GRID;
    if (h->grid_cap > 0 && h->grid_cap < cap) cap = h->grid_cap;
    // contiguous chunk per CTA: a multiple of the block size when the filter is large enough to give
    // every warp full 32-particle steps, else (forced grid mode on a small filter) a multiple of `warps`
    int chunk = (N + cap - 1) / cap;
    if (chunk >= block || h->force_mode != 2) chunk = (chunk + block - 1) / block * block;
    else chunk = (chunk + warps - 1) / warps * warps;
    const int G = (N + chunk - 1) / chunk;
    p.cta_chunk = chunk;
    p.coop = G > 1 ? 1 : 0;
    h->info[0] = G; h->info[1] = 1; h->info[2] = block; h->info[3] = p.coop;
    for (int t = t0; t < t0 + nt; ++t) {
      p.tracker_base = t;
      void* args[] = {(void*)&p};
      dim3 grid((unsigned)G, 1, 1);
      if (p.coop) PF_CUDA(h, cudaLaunchCooperativeKernel((const void*)k, grid, dim3(block), args, smem, h->stream));
      else { k<<<grid, block, smem, h->stream>>>(p); PF_CUDA(h, cudaGetLastError()); }
      h->launches += 1;
    }
  }
  if (split_tail) {
    p.tracker_base = t0;
    tail_kernel_t tk = tail_kernel_for(h->cfg.n_leds);
    tk<<<(unsigned)((nt + PF_TAIL_WARPS - 1) / PF_TAIL_WARPS), 32 * PF_TAIL_WARPS, 0, h->stream>>>(p, frame0);
    PF_CUDA(h, cudaGetLastError());
    h->launches += 1;
  }
  if (timed) PF_CUDA(h, cudaEventRecord(h->ev1, h->stream));
  h->timing_valid = timed;
  return PFMPE_OK;
}

void cm16_to_planes(const double* poses, int N, float* planes /* 3*N*4 */)
{
  for (int n = 0; n < N; ++n) {
    const double* m = poses + 16 * (size_t)n;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 4; ++j) planes[((size_t)i * N + n) * 4 + j] = (float)m[i + 4 * j];
  }
}

void planes_to_cm16(const float* planes, int N, double* poses)
{
  for (int n = 0; n < N; ++n) {
    double* m = poses + 16 * (size_t)n;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 4; ++j) m[i + 4 * j] = (double)planes[((size_t)i * N + n) * 4 + j];
    m[3] = 0; m[7] = 0; m[11] = 0; m[15] = 1;
  }
}

__global__ void pf_fill_kernel(float4* base, int N, int S, float4 r0, float4 r1, float4 r2)
{
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  base[n] = r0;
  base[(size_t)S + n] = r1;
  base[2 * (size_t)S + n] = r2;
}

// pfmpe_track_frame_device_batch: the host finished the Gauss-Newton of n objects (list[], records in pinned host memory,
// read through the unified address space): their refined pose / covariance replace the capped iterate in the device state
__global__ void pf_track_patch_kernel(pf_tracker_state* state, const pf_frame_result* rec, const int* list, int n)
{
  const int q = blockIdx.x, e = threadIdx.x;      // one block per object, 64 threads: 12 pose + 36 covariance entries
  if (q >= n) return;
  const int t = list[q];
  if (e < 12) { const double v = rec[t].pose[e]; state[t].cur[e] = v; state[t].pred[e] = v; }
  else if (e < 48) state[t].cov[e - 12] = rec[t].cov[e - 12];
}

// pfmpe_track_frame_device_batch: the records of the objects whose device Gauss-Newton was capped (gn_iters == -2), packed
// and written straight into pinned host memory (hdr[0] = how many, hdr[4..] = which trackers, out[] = their records), so
// that the host can start on them while the full record array is still crossing the bus.  One CTA; more than
// PF_LATE_MAX objects: hdr[0] = -1, the host then picks them out of the full array.
#define PF_LATE_MAX 2048
__global__ void pf_track_late_kernel(const pf_frame_result* rec, int T, int* hdr, pf_frame_result* out)
{
  static_assert(sizeof(pf_frame_result) % 4 == 0, "records are copied as 32-bit words");
  __shared__ int cnt;
  __shared__ int lst[PF_LATE_MAX];
  if (threadIdx.x == 0) cnt = 0;
  __syncthreads();
  for (int t = threadIdx.x; t < T; t += blockDim.x)
    if (rec[t].status != 0 && rec[t].resampled && rec[t].gn_iters == -2) {
      const int q = atomicAdd(&cnt, 1);
      if (q < PF_LATE_MAX) lst[q] = t;
    }
  __syncthreads();
  const int n = cnt;
  if (n > PF_LATE_MAX) { if (threadIdx.x == 0) hdr[0] = -1; return; }
  if (threadIdx.x == 0) hdr[0] = n;
  for (int q = threadIdx.x; q < n; q += blockDim.x) hdr[4 + q] = lst[q];
  const int W = sizeof(pf_frame_result) / 4;
  for (int q = threadIdx.x >> 5; q < n; q += blockDim.x >> 5) {
    const uint32_t* a = reinterpret_cast<const uint32_t*>(rec + lst[q]);
    uint32_t* b = reinterpret_cast<uint32_t*>(out + q);
    for (int i = threadIdx.x & 31; i < W; i += 32) b[i] = a[i];
  }
}

// track mode: stage the constants of `frame` from the device-resident tracker state
__global__ void pf_track_prepare_kernel(pf_kparams p, int frame)
{
  const int trk = blockIdx.x * blockDim.x + threadIdx.x;
  if (trk >= p.n_trackers) return;
  pf_tracker_state* st = p.state + trk;
  if (st->failed) return;
  pf_aff cur, prev, P, pred;
  for (int i = 0; i < 12; ++i) { cur.a[i] = st->cur[i]; prev.a[i] = st->prev[i]; }
  const double t_next = p.times[frame];
  pf_predict_pose(prev, cur, st->t_prev, st->t_cur, t_next, P, pred);
  st->t_pred = t_next;
  for (int i = 0; i < 12; ++i) st->pred[i] = pred.a[i];
  const size_t fi = (size_t)trk * p.total_frames + frame;
  pf_make_frame_const(p.fconst[trk], cur, pred, P, pf_aff_identity(), st->isi, t_next - st->t_cur, p.seeds[fi],
                      p.blob_count[fi], p.min_ang, p.max_ang, p.min_tr, p.max_tr);
}

}  // namespace

// =========================================================================================
// C ABI
// =========================================================================================
extern "C" {

int pfmpe_version(void) { return PFMPE_VERSION; }

const char* pfmpe_strerror(int s)
{
  switch (s) {
    case PFMPE_OK: return "ok";
    case PFMPE_ERR_INVALID: return "invalid argument";
    case PFMPE_ERR_CUDA: return "CUDA error";
    case PFMPE_ERR_NO_DEVICE: return "no usable CUDA device (no CPU fallback exists)";
    case PFMPE_ERR_STATE: return "call sequence error";
    case PFMPE_ERR_UNSUPPORTED: return "unsupported configuration";
    case PFMPE_ERR_COMM: return "communicator error";
    default: return "unknown status";
  }
}

const char* pfmpe_last_error(const pfmpe_handle* h) { return h ? h->err.c_str() : g_create_err.c_str(); }

int pfmpe_create(const pfmpe_config* cfg, pfmpe_handle** out)
{
  if (!cfg || !out) return PFMPE_ERR_INVALID;
  *out = nullptr;
  if (cfg->n_particles < 2 || cfg->n_trackers < 1 || cfg->n_leds < 1 || cfg->n_leds > PFMPE_MAX_LED ||
      cfg->max_blobs < 1 || cfg->max_blobs > 2048) {
    g_create_err = "pfmpe_create: n_particles >= 2, n_trackers >= 1, 1 <= n_leds <= 8, 1 <= max_blobs <= 2048";
    return PFMPE_ERR_INVALID;
  }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || cfg->device < 0 || cfg->device >= ndev) {
    g_create_err = "pfmpe_create: no usable CUDA device";
    (void)cudaGetLastError();
    return PFMPE_ERR_NO_DEVICE;
  }
  pfmpe_handle* h = new (std::nothrow) pfmpe_handle();
  if (!h) return PFMPE_ERR_INVALID;
  h->cfg = *cfg;
  h->force_mode = cfg->reserved[0];
  h->grid_cap = cfg->reserved[1];
  const size_t N = (size_t)cfg->n_particles, T = (size_t)cfg->n_trackers;
  h->particles_set.assign(T, 0);
  h->markers.assign(T * PF_MAX_LED * 3, 0.0);
  h->last_sum_fx.assign(T, 0);
  h->host_trk.assign(T, HostTrack());
  h->in_scratch.resize(T);
  h->out_scratch.resize(T);
#define PF_TRY(call)                                                              \
  do {                                                                            \
    cudaError_t e_ = (call);                                                      \
    if (e_ != cudaSuccess) {                                                      \
      g_create_err = std::string(#call) + ": " + cudaGetErrorString(e_);          \
      free_all(h); delete h;                                                      \
      return PFMPE_ERR_CUDA;                                                      \
    }                                                                             \
  } while (0)
  PF_TRY(cudaSetDevice(cfg->device));
  cudaDeviceProp prop;
  PF_TRY(cudaGetDeviceProperties(&prop, cfg->device));
  h->num_sms = prop.multiProcessorCount;
  PF_TRY(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  h->own_stream = true;
  PF_TRY(cudaEventCreate(&h->ev0));
  PF_TRY(cudaEventCreate(&h->ev1));
  h->pstride = N;
  PF_TRY(dev_alloc(&h->pose0, T * 3 * N));
  PF_TRY(dev_alloc(&h->pose1, T * 3 * N));
  PF_TRY(dev_alloc(&h->gather_tmp, 3 * N));
  PF_TRY(dev_alloc(&h->w, T * N));
  PF_TRY(dev_alloc(&h->kend, T * N));
  PF_TRY(dev_alloc(&h->idx, T * N));
  PF_TRY(dev_alloc(&h->gather_idx, N));
  PF_TRY(dev_alloc(&h->cta_agg, (size_t)PF_MAX_GRID));
  PF_TRY(dev_alloc(&h->accum, 4));
  h->gfill_cap = (int)(N / 128 + 4096);   // > N / PF_FILL_STEP + N / PF_FILL_SPLIT pieces
  PF_TRY(cudaMalloc((void**)&h->gfill, sizeof(pf_gfill) + (size_t)h->gfill_cap * 16));
  PF_TRY(cudaMemset(h->gfill, 0, sizeof(pf_gfill)));
  h->blobs_off = (T * sizeof(pf_frame_const) + 255) / 256 * 256;
  h->inbox_bytes = h->blobs_off + T * (size_t)cfg->max_blobs * 2 * sizeof(double);
  PF_TRY(cudaMalloc((void**)&h->inbox, h->inbox_bytes));
  h->fconst = reinterpret_cast<pf_frame_const*>(h->inbox);
  h->blobs = reinterpret_cast<double*>(h->inbox + h->blobs_off);
  PF_TRY(dev_alloc(&h->state, T));
  PF_TRY(dev_alloc(&h->result, T));
  PF_TRY(dev_alloc(&h->obj_f, T * PF_MAX_LED * 3));
  PF_TRY(dev_alloc(&h->obj_d, T * PF_MAX_LED * 3));
  PF_TRY(dev_alloc(&h->rng_tab, PF_RNG_TAB_WORDS));
  PF_TRY(dev_alloc(&h->roi_keys, T * 4));
  PF_TRY(dev_alloc(&h->roi_out, T * 4));
  PF_TRY(cudaMallocHost((void**)&h->h_roi, T * 4 * sizeof(float)));
  if (cfg->flags & PFMPE_FLAG_MEAN_COV) {
    const size_t nb = (N + PF_MC_THREADS - 1) / PF_MC_THREADS, lim = T == 1 ? PF_MC_MAXBLK : 16;
    h->mc_blocks = (int)(nb < lim ? nb : lim);
    PF_TRY(dev_alloc(&h->mc_part, T * (size_t)h->mc_blocks * 21));
    PF_TRY(dev_alloc(&h->mc_mean, T * 12));
  }
  PF_TRY(dev_alloc(&h->timeline, 16));
  PF_TRY(cudaMemset(h->timeline, 0, 16 * sizeof(unsigned long long)));
  PF_TRY(cudaMallocHost((void**)&h->h_inbox, h->inbox_bytes));
  PF_TRY(cudaMallocHost((void**)&h->h_result, T * sizeof(pf_frame_result)));
  PF_TRY(cudaMallocHost((void**)&h->h_state, T * sizeof(pf_tracker_state)));
  std::memset(h->h_inbox, 0, h->inbox_bytes);
  PF_TRY(cudaMemset(h->cta_agg, 0, (size_t)PF_MAX_GRID * sizeof(pf_cta_agg)));
  PF_TRY(cudaMemset(h->accum, 0, 4 * sizeof(pf_accum)));
  PF_TRY(cudaMemset(h->w, 0, T * N * sizeof(float)));
  PF_TRY(cudaMemset(h->kend, 0, T * N * sizeof(int)));
  PF_TRY(cudaMemset(h->idx, 0, T * N * sizeof(int)));
  PF_TRY(cudaMemset(h->result, 0, T * sizeof(pf_frame_result)));
  std::memset(h->h_state, 0, T * sizeof(pf_tracker_state));
  PF_TRY(cudaMemcpy(h->state, h->h_state, T * sizeof(pf_tracker_state), cudaMemcpyHostToDevice));
  {
    std::vector<uint32_t> tab(PF_RNG_TAB_WORDS);
    pf_build_rng_table(tab.data());
    if (pf_rng_skip(1u, 384, tab.data()) != PF_RNG_A384) {
      g_create_err = "pfmpe_create: RNG stride constant mismatch";
      free_all(h); delete h;
      return PFMPE_ERR_INVALID;
    }
    PF_TRY(cudaMemcpy(h->rng_tab, tab.data(), PF_RNG_TAB_WORDS * sizeof(uint32_t), cudaMemcpyHostToDevice));
  }
#undef PF_TRY
  *out = h;
  return PFMPE_OK;
}

int pfmpe_destroy(pfmpe_handle* h)
{
  if (!h) return PFMPE_ERR_INVALID;
  cudaSetDevice(h->cfg.device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  free_all(h);
  delete h;
  return PFMPE_OK;
}

int pfmpe_set_markers(pfmpe_handle* h, int tracker, const double* xyz, int n_leds)
{
  if (!h || !xyz) return PFMPE_ERR_INVALID;
  if (n_leds != h->cfg.n_leds) return fail(h, PFMPE_ERR_INVALID, "n_leds differs from pfmpe_config.n_leds");
  if (tracker < -1 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  const int T = h->cfg.n_trackers;
  for (int t = 0; t < T; ++t)
    if (tracker < 0 || tracker == t) std::memcpy(&h->markers[(size_t)t * PF_MAX_LED * 3], xyz, sizeof(double) * 3 * n_leds);
  std::vector<float> f(h->markers.size());
  for (size_t i = 0; i < f.size(); ++i) f[i] = (float)h->markers[i];
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  PF_CUDA(h, cudaMemcpy(h->obj_d, h->markers.data(), h->markers.size() * sizeof(double), cudaMemcpyHostToDevice));
  PF_CUDA(h, cudaMemcpy(h->obj_f, f.data(), f.size() * sizeof(float), cudaMemcpyHostToDevice));
  h->have_markers = true;
  return PFMPE_OK;
}

int pfmpe_set_camera(pfmpe_handle* h, const double K[9], const double* D5)
{
  // the reference's PF projects with K only (project2d); D is used by PFMPE_FLAG_DISTORT (row f3)
  if (!h || !K) return PFMPE_ERR_INVALID;
  std::memcpy(h->K, K, sizeof h->K);
  for (int i = 0; i < 5; ++i) h->D5[i] = D5 ? D5[i] : 0.0;
  h->have_camera = true;
  return PFMPE_OK;
}

int pfmpe_set_params(pfmpe_handle* h, const pfmpe_pf_params* prm)
{
  if (!h || !prm) return PFMPE_ERR_INVALID;
  if (!(prm->back_projection_pixel_tolerance > 0)) return fail(h, PFMPE_ERR_INVALID, "back_projection_pixel_tolerance must be > 0");
  h->prm = *prm;
  h->have_params = true;
  return PFMPE_OK;
}

int pfmpe_set_particles(pfmpe_handle* h, int tracker, const double* poses, int n)
{
  if (!h || !poses) return PFMPE_ERR_INVALID;
  if (n != h->cfg.n_particles) return fail(h, PFMPE_ERR_INVALID, "n differs from pfmpe_config.n_particles");
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  const size_t N = (size_t)n;
  std::vector<float> planes(3 * N * 4);
  cm16_to_planes(poses, n, planes.data());
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  const size_t S = h->pstride;
  for (int pl = 0; pl < 3; ++pl) {
    PF_CUDA(h, cudaMemcpy(h->pose0 + (size_t)tracker * 3 * S + pl * S, planes.data() + (size_t)pl * N * 4, N * sizeof(float4), cudaMemcpyHostToDevice));
    PF_CUDA(h, cudaMemcpy(h->pose1 + (size_t)tracker * 3 * S + pl * S, planes.data() + (size_t)pl * N * 4, N * sizeof(float4), cudaMemcpyHostToDevice));
  }
  h->sh_src = 0;
  PF_CUDA(h, cudaMemcpy(h->h_state + tracker, h->state + tracker, sizeof(pf_tracker_state), cudaMemcpyDeviceToHost));
  pf_tracker_state& s = h->h_state[tracker];
  s.src_buf = 0; s.src_mode = 0; s.t_buf = 1; s.failed = 0;
  PF_CUDA(h, cudaMemcpy(h->state + tracker, &s, sizeof s, cudaMemcpyHostToDevice));
  h->particles_set[tracker] = 1;
  return PFMPE_OK;
}

static int check_ready(pfmpe_handle* h)
{
  if (!h->have_markers || !h->have_camera || !h->have_params)
    return fail(h, PFMPE_ERR_STATE, "set_markers, set_camera and set_params must be called first");
  return PFMPE_OK;
}

static int check_frame(pfmpe_handle* h, int t, const pfmpe_frame_in* in)
{
  if (!h->particles_set[t]) return fail(h, PFMPE_ERR_STATE, "set_particles / track_init must precede update");
  if (in->n_image_points < 0 || in->n_image_points > h->cfg.max_blobs) return fail(h, PFMPE_ERR_INVALID, "n_image_points exceeds max_blobs");
  if (in->n_image_points > 0 && !in->image_points) return fail(h, PFMPE_ERR_INVALID, "image_points is NULL");
  return PFMPE_OK;
}

static int stage_frame(pfmpe_handle* h, int t, const pfmpe_frame_in* in)
{
  pf_frame_const* fc = reinterpret_cast<pf_frame_const*>(h->h_inbox) + t;
  pf_make_frame_const(*fc, pf_aff_from_cm16(in->current_pose), pf_aff_from_cm16(in->predicted_pose),
                      pf_aff_from_cm16(in->prediction_matrix), pf_aff_from_cm16(in->cam_move_inv),
                      in->it_since_initialized, in->dt, in->seed, in->n_image_points, h->prm.min_angular_noise,
                      h->prm.max_angular_noise, h->prm.min_transition_noise, h->prm.max_transition_noise);
  double* b = reinterpret_cast<double*>(h->h_inbox + h->blobs_off) + (size_t)t * h->cfg.max_blobs * 2;
  if (in->n_image_points) std::memcpy(b, in->image_points, sizeof(double) * 2 * (size_t)in->n_image_points);
  return PFMPE_OK;
}

static void unpack_result(const pf_frame_result& r, pfmpe_frame_out* o)
{
  std::memset(o, 0, sizeof *o);
  o->status = r.status;
  o->iterations = r.iterations;
  o->best_iteration = r.best_iteration;
  o->resampled = r.resampled;
  o->highest_prob = (double)r.highest_prob;
  o->weight_sum = (double)r.sum_fx / 1073741824.0;
  o->most_likely_particle_idx = r.most_likely_particle_idx;
  o->most_likely_pose_idx = r.most_likely_pose_idx;
  o->max_count = r.max_count;
  o->n_correspondences = r.npairs;
  for (int i = 0; i < 2 * PFMPE_MAX_LED; ++i) o->correspondences[i] = r.pairs[i];
  pf_aff T;
  for (int i = 0; i < 12; ++i) T.a[i] = (double)r.winner_pose[i];
  pf_aff_to_cm16(T, o->winner_pose);
  for (int i = 0; i < 12; ++i) T.a[i] = r.mean_pose[i];
  pf_aff_to_cm16(T, o->mean_pose);
  for (int i = 0; i < 36; ++i) o->mean_cov[i] = r.mean_cov[i];
}

static int run_update(pfmpe_handle* h, int t0, int nt, const pfmpe_frame_in* in, pfmpe_frame_out* out)
{
  int rc = check_ready(h);
  if (rc) return rc;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  for (int i = 0; i < nt; ++i) {
    rc = check_frame(h, t0 + i, in + i);
    if (rc) return rc;
  }
  host_parallel_for(h, nt, [&](int i) { stage_frame(h, t0 + i, in + i); });
  const int T = h->cfg.n_trackers;
  const bool want_mean = (h->cfg.flags & PFMPE_FLAG_MEAN_COV) != 0;
  // latency path: one tracker with few detections carries its inputs in the kernel parameters; the
  // result record is written by the kernel straight into pinned host memory (no copies either way)
  InlineArgs ia;
  const bool inline_in = nt == 1 && in[0].n_image_points <= PF_INLINE_BLOBS;
  if (inline_in) {
    ia.fc = reinterpret_cast<const pf_frame_const*>(h->h_inbox) + t0;
    ia.blobs = reinterpret_cast<const double*>(h->h_inbox + h->blobs_off) + (size_t)t0 * h->cfg.max_blobs * 2;
    ia.D = in[0].n_image_points;
  } else if (nt == T) {
    PF_CUDA(h, cudaMemcpyAsync(h->inbox, h->h_inbox, h->inbox_bytes, cudaMemcpyHostToDevice, h->stream));
  } else {
    PF_CUDA(h, cudaMemcpyAsync(h->fconst + t0, reinterpret_cast<pf_frame_const*>(h->h_inbox) + t0,
                               sizeof(pf_frame_const) * nt, cudaMemcpyHostToDevice, h->stream));
    const size_t off = (size_t)t0 * h->cfg.max_blobs * 2;
    PF_CUDA(h, cudaMemcpyAsync(h->blobs + off, reinterpret_cast<double*>(h->h_inbox + h->blobs_off) + off,
                               sizeof(double) * 2 * (size_t)h->cfg.max_blobs * nt, cudaMemcpyHostToDevice, h->stream));
  }
  if (!want_mean) ia.result = h->h_result;
  const bool poll = !want_mean && nt == 1;
  if (poll) {
    h->done_seq = h->done_seq == 0x7fffffff ? 1 : h->done_seq + 1;
    ia.done_seq = h->done_seq;
    reinterpret_cast<volatile int32_t*>(&h->h_result[t0].pad)[0] = 0;
  }
  rc = launch_frames(h, 0, t0, nt, 0, 1, nullptr, &ia);
  if (rc) return rc;
  if (poll) {
    // the record is complete as soon as its sequence word arrives (written last, after a system-scope
    // fence); the rest of the kernel (state mirror, CTA exit) overlaps with the host tail.  Later calls
    // are ordered behind it on the stream.  cudaStreamQuery guards against a failed launch.
    volatile int32_t* flag = reinterpret_cast<volatile int32_t*>(&h->h_result[t0].pad);
    for (unsigned spin = 1;; ++spin) {
      if (*flag == ia.done_seq) break;
      if ((spin & 0x3fffu) == 0) {
        const cudaError_t q = cudaStreamQuery(h->stream);
        if (q == cudaSuccess) { if (*flag == ia.done_seq) break; return fail(h, PFMPE_ERR_CUDA, "frame kernel retired without a result"); }
        if (q != cudaErrorNotReady) { h->err = std::string("frame kernel: ") + cudaGetErrorString(q); return PFMPE_ERR_CUDA; }
      }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    unpack_result(h->h_result[t0], out);
    h->last_sum_fx[t0] = h->h_result[t0].sum_fx;
    h->last_was_track = false;
    return PFMPE_OK;
  }
  if (want_mean) {
    rc = launch_meancov(h, t0, nt);
    if (rc) return rc;
    PF_CUDA(h, cudaMemcpyAsync(h->h_result + t0, h->result + t0, sizeof(pf_frame_result) * nt, cudaMemcpyDeviceToHost, h->stream));
  }
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  host_parallel_for(h, nt, [&](int i) {
    unpack_result(h->h_result[t0 + i], out + i);
    h->last_sum_fx[t0 + i] = h->h_result[t0 + i].sum_fx;
  });
  h->last_was_track = false;
  return PFMPE_OK;
}

int pfmpe_update(pfmpe_handle* h, int tracker, const pfmpe_frame_in* in, pfmpe_frame_out* out)
{
  if (!h || !in || !out) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  return run_update(h, tracker, 1, in, out);
}

int pfmpe_update_batch(pfmpe_handle* h, const pfmpe_frame_in* in, pfmpe_frame_out* out)
{
  if (!h || !in || !out) return PFMPE_ERR_INVALID;
  return run_update(h, 0, h->cfg.n_trackers, in, out);
}

// ---- getters ------------------------------------------------------------------------------
static int fetch_state(pfmpe_handle* h, int tracker)
{
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  PF_CUDA(h, cudaMemcpy(h->h_state + tracker, h->state + tracker, sizeof(pf_tracker_state), cudaMemcpyDeviceToHost));
  return PFMPE_OK;
}

// run the gather kernel: gather_tmp = newPoseEstimation, gather_idx = Particle_index per slot
static int run_gather(pfmpe_handle* h, int tracker)
{
  const pf_tracker_state& s = h->h_state[tracker];
  const size_t N = (size_t)h->cfg.n_particles;
  const float4* T = (s.src_buf == 0 ? h->pose0 : h->pose1) + (size_t)tracker * 3 * h->pstride;
  const int blocks = (int)((N + PF_BLOCK - 1) / PF_BLOCK);
  pf_gather_kernel<<<blocks, PF_BLOCK, 0, h->stream>>>(T, (h->sharded && h->lead) ? h->lead : h->idx + (size_t)tracker * N, (int)N, (int)h->pstride, s.src_mode,
                                                      h->gather_tmp, h->gather_idx);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  return PFMPE_OK;
}

int pfmpe_get_particles(pfmpe_handle* h, int tracker, int which, double* poses)
{
  if (!h || !poses) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  int rc = fetch_state(h, tracker);
  if (rc) return rc;
  const size_t N = (size_t)h->cfg.n_particles;
  std::vector<float> planes(3 * N * 4);
  const float4* src;
  size_t S = N;
  if (which == 0) {
    S = h->pstride;
    src = (h->h_state[tracker].t_buf == 0 ? h->pose0 : h->pose1) + (size_t)tracker * 3 * S;
  } else {
    rc = run_gather(h, tracker);
    if (rc) return rc;
    src = h->gather_tmp;
  }
  for (int pl = 0; pl < 3; ++pl)
    PF_CUDA(h, cudaMemcpy(planes.data() + (size_t)pl * N * 4, src + pl * S, N * sizeof(float4), cudaMemcpyDeviceToHost));
  planes_to_cm16(planes.data(), (int)N, poses);
  return PFMPE_OK;
}

static int current_sum_fx(pfmpe_handle* h, int tracker, long long* s)
{
  if (h->last_was_track && h->seq.last_frame >= 0) {
    pf_frame_result r;
    PF_CUDA(h, cudaMemcpy(&r, h->seq.result + (size_t)tracker * h->seq.n_frames + h->seq.last_frame, sizeof r,
                          cudaMemcpyDeviceToHost));
    *s = r.sum_fx;
  } else *s = h->last_sum_fx[tracker];
  return PFMPE_OK;
}

int pfmpe_get_weights(pfmpe_handle* h, int tracker, int normalised, double* w)
{
  if (!h || !w) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  const size_t N = (size_t)h->cfg.n_particles;
  std::vector<float> f(N);
  PF_CUDA(h, cudaMemcpy(f.data(), h->w + (size_t)tracker * N, N * sizeof(float), cudaMemcpyDeviceToHost));
  double S = 1.0;
  if (normalised) {
    long long sfx = 0;
    int rc = current_sum_fx(h, tracker, &sfx);
    if (rc) return rc;
    S = sfx != 0 ? (double)sfx / 1073741824.0 : 1.0;   // probPart / probPartSum only if the sum is non-zero (:628)
  }
  for (size_t i = 0; i < N; ++i) w[i] = (double)f[i] / S;
  return PFMPE_OK;
}

int pfmpe_get_resample_indices(pfmpe_handle* h, int tracker, int32_t* idx)
{
  if (!h || !idx) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  int rc = fetch_state(h, tracker);
  if (rc) return rc;
  if (h->h_state[tracker].src_mode != 1) return fail(h, PFMPE_ERR_STATE, "the last frame of this tracker was not resampled");
  rc = run_gather(h, tracker);
  if (rc) return rc;
  PF_CUDA(h, cudaMemcpy(idx, h->gather_idx, (size_t)h->cfg.n_particles * sizeof(int), cudaMemcpyDeviceToHost));
  return PFMPE_OK;
}

int pfmpe_get_resample_counts(pfmpe_handle* h, int tracker, uint32_t* count)
{
  if (!h || !count) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  int rc = fetch_state(h, tracker);
  if (rc) return rc;
  if (h->h_state[tracker].src_mode != 1) return fail(h, PFMPE_ERR_STATE, "the last frame of this tracker was not resampled");
  const size_t N = (size_t)h->cfg.n_particles;
  std::vector<int> ke(N);
  PF_CUDA(h, cudaMemcpy(ke.data(), h->kend + (size_t)tracker * N, N * sizeof(int), cudaMemcpyDeviceToHost));
  int prev = 0;
  for (size_t i = 0; i < N; ++i) { count[i] = (uint32_t)(ke[i] - prev); prev = ke[i]; }
  return PFMPE_OK;
}

// ---- host tail ----------------------------------------------------------------------------
int pfmpe_optimise_pose(const double K[9], const double* markers_xyz, int n_leds, const double* image_points,
                        int n_image_points, const uint32_t* correspondences, int n_correspondences,
                        double pose[16], double cov[36], int* iters)
{
  if (!K || !markers_xyz || !pose || !cov || n_correspondences < 0 || n_image_points < 0 ||
      (n_correspondences > 0 && (!correspondences || !image_points || n_image_points == 0)))
    return PFMPE_ERR_INVALID;
  for (int i = 0; i < n_correspondences; ++i) {   // pairs are 1-based: LED in 1..n_leds, detection in 1..n_image_points
    if (correspondences[2 * i] < 1 || correspondences[2 * i] > (unsigned)n_leds) return PFMPE_ERR_INVALID;
    if (correspondences[2 * i + 1] < 1 || correspondences[2 * i + 1] > (unsigned)n_image_points) return PFMPE_ERR_INVALID;
  }
  pf_aff T = pf_aff_from_cm16(pose);
  double c[36];
  const int it = pf_optimise_pose(K, markers_xyz, image_points, correspondences, n_correspondences, T, c);
  pf_aff_to_cm16(T, pose);
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) cov[i + 6 * j] = c[6 * i + j];
  if (iters) *iters = it;
  return PFMPE_OK;
}

int pfmpe_predict_pose(const double previous_pose[16], const double current_pose[16], double previous_time,
                       double current_time, double time_to_predict, double prediction_matrix[16],
                       double predicted_pose[16])
{
  if (!previous_pose || !current_pose || !prediction_matrix || !predicted_pose) return PFMPE_ERR_INVALID;
  pf_aff P, pred;
  pf_predict_pose(pf_aff_from_cm16(previous_pose), pf_aff_from_cm16(current_pose), previous_time, current_time,
                  time_to_predict, P, pred);
  pf_aff_to_cm16(P, prediction_matrix);
  pf_aff_to_cm16(pred, predicted_pose);
  return PFMPE_OK;
}

int pfmpe_exponential_map(const double twist[6], double pose[16])
{
  if (!twist || !pose) return PFMPE_ERR_INVALID;
  pf_aff_to_cm16(pf_exp_map(twist), pose);
  return PFMPE_OK;
}

int pfmpe_logarithm_map(const double pose[16], double twist[6])
{
  if (!pose || !twist) return PFMPE_ERR_INVALID;
  pf_log_map(pf_aff_from_cm16(pose), twist);
  return PFMPE_OK;
}

// ---- device-resident tracking ----------------------------------------------------------------
int pfmpe_track_init(pfmpe_handle* h, int tracker, const double pose0[16], double t0, const double pose1[16], double t1)
{
  if (!h || !pose0 || !pose1) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  if (!(t1 != t0)) return fail(h, PFMPE_ERR_INVALID, "t0 and t1 must differ (predictPose divides by their difference, :1002)");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  pf_tracker_state s;
  PF_CUDA(h, cudaMemcpyAsync(&h->h_state[tracker], h->state + tracker, sizeof s, cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  s = h->h_state[tracker];
  const pf_aff a0 = pf_aff_from_cm16(pose0), a1 = pf_aff_from_cm16(pose1);
  {
    HostTrack& ht = h->host_trk[tracker];
    ht.prev = a0; ht.cur = a1; ht.t_prev = t0; ht.t_cur = t1; ht.isi = 2; ht.valid = true;
    for (int i = 0; i < 36; ++i) ht.cov[i] = 0.0;
  }
  for (int i = 0; i < 12; ++i) { s.prev[i] = a0.a[i]; s.cur[i] = a1.a[i]; s.pred[i] = a1.a[i]; }
  s.t_prev = t0; s.t_cur = t1; s.t_pred = t1;
  for (int i = 0; i < 36; ++i) s.cov[i] = 0.0;
  s.isi = 2; s.failed = 0; s.src_buf = 0; s.src_mode = 0; s.t_buf = 1; s.frames_done = 0;
  h->h_state[tracker] = s;
  PF_CUDA(h, cudaMemcpyAsync(h->state + tracker, &h->h_state[tracker], sizeof s, cudaMemcpyHostToDevice, h->stream));
  const pf_pose f = pf_aff_to_f32(a1);
  const int N = h->cfg.n_particles;
  const float4 r0 = make_float4(f.p[0], f.p[1], f.p[2], f.p[3]), r1 = make_float4(f.p[4], f.p[5], f.p[6], f.p[7]),
               r2 = make_float4(f.p[8], f.p[9], f.p[10], f.p[11]);
  const int blocks = (N + 255) / 256;
  const int S = (int)h->pstride;
  pf_fill_kernel<<<blocks, 256, 0, h->stream>>>(h->pose0 + (size_t)tracker * 3 * S, N, S, r0, r1, r2);
  pf_fill_kernel<<<blocks, 256, 0, h->stream>>>(h->pose1 + (size_t)tracker * 3 * S, N, S, r0, r1, r2);
  h->sh_src = 0;
  PF_CUDA(h, cudaGetLastError());
  h->launches += 2;
  PF_CUDA(h, cudaStreamSynchronize(h->stream));   // h_state is reused by the next call
  h->particles_set[tracker] = 1;
  h->seq.next_frame = -1;
  return PFMPE_OK;
}

int pfmpe_track_set_state(pfmpe_handle* h, int tracker, const double previous_pose[16], double previous_time,
                          const double current_pose[16], double current_time, uint32_t it_since_initialized)
{
  if (!h || !previous_pose || !current_pose) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  if (!h->particles_set[tracker]) return fail(h, PFMPE_ERR_STATE, "set_particles must precede track_set_state");
  // predictPose divides by (current_time - previous_time) (:1002): equal times would hand inf / NaN constants to the kernel
  if (!(current_time != previous_time)) return fail(h, PFMPE_ERR_INVALID, "previous_time and current_time must differ");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  PF_CUDA(h, cudaMemcpy(h->h_state + tracker, h->state + tracker, sizeof(pf_tracker_state), cudaMemcpyDeviceToHost));
  pf_tracker_state& s = h->h_state[tracker];
  const pf_aff a0 = pf_aff_from_cm16(previous_pose), a1 = pf_aff_from_cm16(current_pose);
  HostTrack& ht = h->host_trk[tracker];
  ht.prev = a0; ht.cur = a1; ht.t_prev = previous_time; ht.t_cur = current_time; ht.isi = it_since_initialized; ht.valid = true;
  for (int i = 0; i < 36; ++i) ht.cov[i] = 0.0;
  for (int i = 0; i < 12; ++i) { s.prev[i] = a0.a[i]; s.cur[i] = a1.a[i]; s.pred[i] = a1.a[i]; }
  s.t_prev = previous_time; s.t_cur = current_time; s.t_pred = current_time;
  for (int i = 0; i < 36; ++i) s.cov[i] = 0.0;
  s.isi = it_since_initialized; s.failed = 0; s.frames_done = 0;
  PF_CUDA(h, cudaMemcpy(h->state + tracker, &s, sizeof s, cudaMemcpyHostToDevice));
  h->seq.next_frame = -1;
  return PFMPE_OK;
}

static int seq_release(pfmpe_handle* h)
{
  if (h->seq.owned) { cudaFree(h->seq.blobs); cudaFree(h->seq.counts); cudaFree(h->seq.times); cudaFree(h->seq.seeds); }
  cudaFree(h->seq.result); cudaFree(h->seq.fc_log);
  h->seq = TrackSeq();
  return PFMPE_OK;
}

static int seq_alloc_results(pfmpe_handle* h, int n_frames)
{
  const size_t n = (size_t)h->cfg.n_trackers * n_frames;
  PF_CUDA(h, dev_alloc(&h->seq.result, n));
  PF_CUDA(h, cudaMemset(h->seq.result, 0, n * sizeof(pf_frame_result)));
  PF_CUDA(h, dev_alloc(&h->seq.fc_log, n));
  PF_CUDA(h, cudaMemset(h->seq.fc_log, 0, n * sizeof(pf_frame_const)));
  return PFMPE_OK;
}

int pfmpe_track_upload(pfmpe_handle* h, int n_frames, const double* image_points, const int32_t* counts,
                       const double* times, const uint32_t* seeds)
{
  if (!h || n_frames < 1 || !image_points || !counts || !times || !seeds) return PFMPE_ERR_INVALID;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  const size_t T = (size_t)h->cfg.n_trackers, F = (size_t)n_frames, D = (size_t)h->cfg.max_blobs;
  for (size_t i = 0; i < T * F; ++i)
    if (counts[i] < 0 || counts[i] > h->cfg.max_blobs) return fail(h, PFMPE_ERR_INVALID, "blob count exceeds max_blobs");
  seq_release(h);
  h->seq.owned = true;
  h->seq.n_frames = n_frames;
  PF_CUDA(h, dev_alloc(&h->seq.blobs, T * F * D * 2));
  PF_CUDA(h, dev_alloc(&h->seq.counts, T * F));
  PF_CUDA(h, dev_alloc(&h->seq.times, F));
  PF_CUDA(h, dev_alloc(&h->seq.seeds, T * F));
  PF_CUDA(h, cudaMemcpy(h->seq.blobs, image_points, T * F * D * 2 * sizeof(double), cudaMemcpyHostToDevice));
  PF_CUDA(h, cudaMemcpy(h->seq.counts, counts, T * F * sizeof(int), cudaMemcpyHostToDevice));
  PF_CUDA(h, cudaMemcpy(h->seq.times, times, F * sizeof(double), cudaMemcpyHostToDevice));
  PF_CUDA(h, cudaMemcpy(h->seq.seeds, seeds, T * F * sizeof(uint32_t), cudaMemcpyHostToDevice));
  return seq_alloc_results(h, n_frames);
}

int pfmpe_track_attach(pfmpe_handle* h, int n_frames, const void* d_image_points, const void* d_counts,
                       const void* d_times, const void* d_seeds)
{
  if (!h || n_frames < 1 || !d_image_points || !d_counts || !d_times || !d_seeds) return PFMPE_ERR_INVALID;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  seq_release(h);
  h->seq.owned = false;
  h->seq.n_frames = n_frames;
  h->seq.blobs = (double*)d_image_points;
  h->seq.counts = (int*)d_counts;
  h->seq.times = (double*)d_times;
  h->seq.seeds = (uint32_t*)d_seeds;
  return seq_alloc_results(h, n_frames);
}

int pfmpe_track_run(pfmpe_handle* h, int first_frame, int n_frames)
{
  if (!h) return PFMPE_ERR_INVALID;
  int rc = check_ready(h);
  if (rc) return rc;
  if (!h->seq.blobs) return fail(h, PFMPE_ERR_STATE, "track_upload / track_attach must precede track_run");
  if (first_frame < 0 || n_frames < 1 || first_frame + n_frames > h->seq.n_frames) return fail(h, PFMPE_ERR_INVALID, "frame range outside the uploaded sequence");
  for (int t = 0; t < h->cfg.n_trackers; ++t)
    if (!h->particles_set[t]) return fail(h, PFMPE_ERR_STATE, "track_init must be called for every tracker");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  if (h->seq.next_frame != first_frame) {
    pf_kparams p;
    fill_kparams(h, p);
    p.blob_count = h->seq.counts; p.times = h->seq.times; p.seeds = h->seq.seeds; p.total_frames = h->seq.n_frames;
    const int T = h->cfg.n_trackers;
    pf_track_prepare_kernel<<<(T + 127) / 128, 128, 0, h->stream>>>(p, first_frame);
    PF_CUDA(h, cudaGetLastError());
    h->launches += 1;
  }
  rc = launch_frames(h, 1, 0, h->cfg.n_trackers, first_frame, n_frames);
  if (rc) return rc;
  h->seq.next_frame = first_frame + n_frames;   // the tail staged the constants of the following frame
  h->seq.last_frame = first_frame + n_frames - 1;
  if (h->seq.last_frame > h->seq.tracked_upto) h->seq.tracked_upto = h->seq.last_frame;
  h->last_was_track = true;
  return PFMPE_OK;
}

int pfmpe_track_replay(pfmpe_handle* h, int first_frame, int n_frames)
{
  if (!h) return PFMPE_ERR_INVALID;
  int rc = check_ready(h);
  if (rc) return rc;
  if (!h->seq.blobs || h->seq.tracked_upto < 0) return fail(h, PFMPE_ERR_STATE, "track_run must have processed the frames before they can be replayed");
  if (first_frame < 0 || n_frames < 1 || first_frame + n_frames > h->seq.tracked_upto + 1) return fail(h, PFMPE_ERR_INVALID, "frame range was not tracked yet");
  for (int t = 0; t < h->cfg.n_trackers; ++t)
    if (!h->particles_set[t]) return fail(h, PFMPE_ERR_STATE, "track_init must be called for every tracker");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  rc = launch_frames(h, 2, 0, h->cfg.n_trackers, first_frame, n_frames);
  if (rc) return rc;
  h->seq.next_frame = -1;                        // fconst no longer holds the constants of a tracked frame
  h->seq.last_frame = first_frame + n_frames - 1;
  h->last_was_track = true;
  return PFMPE_OK;
}

int pfmpe_track_results(pfmpe_handle* h, int tracker, int first_frame, int n_frames, pfmpe_track_result* out)
{
  if (!h || !out) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  if (!h->seq.result || first_frame < 0 || n_frames < 1 || first_frame + n_frames > h->seq.n_frames)
    return fail(h, PFMPE_ERR_INVALID, "frame range outside the uploaded sequence");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  std::vector<pf_frame_result> r((size_t)n_frames);
  PF_CUDA(h, cudaMemcpy(r.data(), h->seq.result + (size_t)tracker * h->seq.n_frames + first_frame,
                        r.size() * sizeof(pf_frame_result), cudaMemcpyDeviceToHost));
  for (int f = 0; f < n_frames; ++f) {
    const pf_frame_result& s = r[f];
    pfmpe_track_result* o = out + f;
    std::memset(o, 0, sizeof *o);
    o->status = s.status; o->flag_fail = s.flag_fail; o->iterations = s.iterations; o->gn_iterations = s.gn_iters;
    o->most_likely_pose_idx = s.most_likely_pose_idx; o->max_count = s.max_count;
    o->highest_prob = s.highest_prob; o->weight_sum = (double)s.sum_fx / 1073741824.0;
    pf_aff T;
    for (int i = 0; i < 12; ++i) T.a[i] = s.pose[i];
    pf_aff_to_cm16(T, o->pose);
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) o->cov[i + 6 * j] = s.cov[6 * i + j];
    for (int i = 0; i < 12; ++i) T.a[i] = (double)s.winner_pose[i];
    pf_aff_to_cm16(T, o->winner_pose);
    o->n_correspondences = s.npairs;
    for (int i = 0; i < 2 * PFMPE_MAX_LED; ++i) o->correspondences[i] = s.pairs[i];
  }
  return PFMPE_OK;
}

// ---- per-frame tracking with HOST buffers: the PF branch of estimateBodyPose (:223-733) minus detection/ROI ----
// tracker i reads image_points + i * stride (doubles), counts[i * cstride], seeds[i * cstride], writes out[i * ostride]
// The FP64 host tail of a batch (predictPose before, Gauss-Newton after the PF update) is independent per tracked
// object: above a few dozen objects it is spread over the host cores (the reference's `for objectNumber` loop, :89, is
// serial; 4096 objects x ~3 us would otherwise dwarf the 1.3 ms the device needs).
}  // extern "C" (C++ helper with internal linkage follows)
static void host_parallel_for(pfmpe_handle* h, int n, const std::function<void(int)>& body)
{
  if (n < 64) { for (int i = 0; i < n; ++i) body(i); return; }   // the single-object latency path never gets here
  if (!h->pool) {
    int nth = (int)std::thread::hardware_concurrency();
    if (const char* e = std::getenv("PFMPE_HOST_THREADS")) nth = std::atoi(e);
    if (nth > 32) nth = 32;
    if (nth < 1) nth = 1;
    h->pool.reset(new HostPool(nth - 1));
  }
  if (h->pool->size() == 0) { for (int i = 0; i < n; ++i) body(i); return; }
  h->pool->run(n, body);
}
extern "C" {

static int track_frames_host(pfmpe_handle* h, int t0, int nt, double time_to_predict, const double* image_points,
                             size_t stride, const int32_t* counts, const uint32_t* seeds, size_t cstride,
                             pfmpe_track_result* out, size_t ostride)
{
  for (int i = 0; i < nt; ++i)
    if (!h->host_trk[t0 + i].valid) return fail(h, PFMPE_ERR_STATE, "track_init must precede track_frame_host");
  host_parallel_for(h, nt, [&](int i) {
    HostTrack& ht = h->host_trk[t0 + i];
    pfmpe_frame_in& in = h->in_scratch[i];
    pf_aff P, pred;
    pf_predict_pose(ht.prev, ht.cur, ht.t_prev, ht.t_cur, time_to_predict, P, pred);   // :234 -> :995-1010
    in.image_points = image_points + (size_t)i * stride;
    in.n_image_points = counts[i * cstride];
    in.it_since_initialized = ht.isi;
    pf_aff_to_cm16(ht.cur, in.current_pose);
    pf_aff_to_cm16(pred, in.predicted_pose);
    pf_aff_to_cm16(P, in.prediction_matrix);
    pf_aff_to_cm16(pf_aff_identity(), in.cam_move_inv);
    in.dt = time_to_predict - ht.t_cur;
    in.seed = seeds[i * cstride];
    in.reserved = 0;
  });
  int rc = run_update(h, t0, nt, h->in_scratch.data(), h->out_scratch.data());
  if (rc) return rc;
  host_parallel_for(h, nt, [&](int i) {
    HostTrack& ht = h->host_trk[t0 + i];
    const pfmpe_frame_out& fo = h->out_scratch[i];
    pfmpe_track_result* o = out + i * ostride;
    std::memset(o, 0, sizeof *o);
    o->status = fo.status; o->flag_fail = (float)fo.status; o->iterations = fo.iterations; o->gn_iterations = -1;
    o->most_likely_pose_idx = fo.most_likely_pose_idx; o->max_count = fo.max_count;
    o->highest_prob = fo.highest_prob; o->weight_sum = fo.weight_sum;
    std::memcpy(o->winner_pose, fo.winner_pose, sizeof o->winner_pose);
    o->n_correspondences = fo.n_correspondences;
    std::memcpy(o->correspondences, fo.correspondences, sizeof o->correspondences);
    if (fo.resampled) {
      pf_aff pose = pf_aff_from_cm16(fo.winner_pose);       // predicted_pose_ = PoseParticle[q] (:687)
      const pf_aff before = pose;
      double c[36];
      o->gn_iterations = pf_optimise_pose(h->K, &h->markers[(size_t)(t0 + i) * PF_MAX_LED * 3], image_points + (size_t)i * stride,
                                          fo.correspondences, fo.n_correspondences, pose, c);   // :690 -> :1805-2009
      if (ht.isi < 2) ++ht.isi;                             // :2027-2030
      ht.prev = ht.cur; ht.cur = pose;                      // updatePose :2013-2014
      if (time_to_predict - ht.t_cur > 0.001 || time_to_predict < ht.t_cur) { ht.t_prev = ht.t_cur; ht.t_cur = time_to_predict; }
      for (int r = 0; r < 3; ++r)
        for (int q = 0; q < 3; ++q)
          if (fabs(before.a[4 * r + q] - pose.a[4 * r + q]) >= 0.3) o->flag_fail = 1.5f;   // :693-701
      for (int r = 0; r < 36; ++r) ht.cov[r] = c[r];
    } else {
      ht.isi = 0;                                           // :710: the caller must re-initialise
    }
    pf_aff_to_cm16(ht.cur, o->pose);
    for (int r = 0; r < 6; ++r) for (int q = 0; q < 6; ++q) o->cov[r + 6 * q] = ht.cov[6 * r + q];
  });
  return PFMPE_OK;
}

int pfmpe_track_frame_host(pfmpe_handle* h, int tracker, double time_to_predict, const double* image_points,
                           int32_t n_image_points, uint32_t seed, pfmpe_track_result* out)
{
  if (!h || !out || (n_image_points > 0 && !image_points)) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  return track_frames_host(h, tracker, 1, time_to_predict, image_points, 0, &n_image_points, &seed, 0, out, 0);
}

int pfmpe_track_frame_host_batch(pfmpe_handle* h, double time_to_predict, const double* image_points,
                                 const int32_t* counts, const uint32_t* seeds, pfmpe_track_result* out)
{
  if (!h || !out || !image_points || !counts || !seeds) return PFMPE_ERR_INVALID;
  return track_frames_host(h, 0, h->cfg.n_trackers, time_to_predict, image_points, (size_t)h->cfg.max_blobs * 2, counts, seeds, 1,
                           out, 1);
}

int pfmpe_track_sequence_host(pfmpe_handle* h, int n_frames, const double* times, const double* image_points,
                              const int32_t* counts, const uint32_t* seeds, pfmpe_track_result* out, double* frame_seconds)
{
  if (!h || n_frames < 1 || !times || !image_points || !counts || !seeds || !out) return PFMPE_ERR_INVALID;
  const int T = h->cfg.n_trackers;
  const size_t D2 = (size_t)h->cfg.max_blobs * 2;
  for (int f = 0; f < n_frames; ++f) {
    const auto a = std::chrono::steady_clock::now();
    const int rc = track_frames_host(h, 0, T, times[f], image_points + (size_t)f * D2, (size_t)n_frames * D2, counts + f, seeds + f,
                                     (size_t)n_frames, out + f, (size_t)n_frames);
    if (rc) return rc;
    if (frame_seconds) frame_seconds[f] = std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count();
  }
  return PFMPE_OK;
}

// ---- per-frame tracking with HOST buffers and the DEVICE tail (row f2) ---------------------------------------------
// The same branch of estimateBodyPose as pfmpe_track_frame_host_batch, but predictPose, the PF update,
// optimiseAndUpdatePose and the jump check all run in ONE launch of the track-mode kernel: the frame's detections go
// host -> device (pinned staging), the result records come back device -> host.  For batches of tracked objects, where
// the FP64 Gauss-Newton of thousands of objects would otherwise run on the host cores.
static int track_frames_device(pfmpe_handle* h, double time_to_predict, const double* image_points, size_t stride,
                               const int32_t* counts, const uint32_t* seeds, size_t cstride, pfmpe_track_result* out, size_t ostride)
{
  int rc = check_ready(h);
  if (rc) return rc;
  const int T = h->cfg.n_trackers, D = h->cfg.max_blobs;
  for (int t = 0; t < T; ++t) {
    if (!h->particles_set[t] || !h->host_trk[t].valid) return fail(h, PFMPE_ERR_STATE, "track_init must be called for every tracker");
    if (counts[t * cstride] < 0 || counts[t * cstride] > D) return fail(h, PFMPE_ERR_INVALID, "blob count exceeds max_blobs");
  }
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  const size_t nb = (size_t)T * D * 2 * sizeof(double), off_time = nb, off_cnt = off_time + 8, off_seed = off_cnt + (size_t)T * 4;
  if (!h->live.blobs) {
    PF_CUDA(h, dev_alloc(&h->live.blobs, (size_t)T * D * 2));
    PF_CUDA(h, dev_alloc(&h->live.counts, (size_t)T));
    PF_CUDA(h, dev_alloc(&h->live.times, (size_t)1));
    PF_CUDA(h, dev_alloc(&h->live.seeds, (size_t)T));
    PF_CUDA(h, dev_alloc(&h->live.result, (size_t)T));
    PF_CUDA(h, cudaMemset(h->live.result, 0, (size_t)T * sizeof(pf_frame_result)));
    PF_CUDA(h, cudaMallocHost((void**)&h->h_live, off_seed + (size_t)T * 4));
    PF_CUDA(h, cudaMallocHost((void**)&h->h_live_res, (size_t)T * sizeof(pf_frame_result)));
    PF_CUDA(h, cudaMallocHost((void**)&h->h_late_hdr, sizeof(int) * (4 + (size_t)T)));
    PF_CUDA(h, cudaMallocHost((void**)&h->h_late_rec, sizeof(pf_frame_result) * (size_t)std::min(T, PF_LATE_MAX)));
    PF_CUDA(h, cudaEventCreateWithFlags(&h->ev_late, cudaEventDisableTiming));
    h->live.n_frames = 1; h->live.owned = true;
  }
  static const bool trace = std::getenv("PFMPE_TRACE") != nullptr;    // per-phase wall times of this call, printed every 20 frames
  static std::vector<std::array<double, 10>> tr;
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t0 = trace ? now() : 0.0;
  double t1 = 0, t2 = 0, t3 = 0, t4 = 0;
  double* sb = reinterpret_cast<double*>(h->h_live);
  int32_t* sc = reinterpret_cast<int32_t*>(h->h_live + off_cnt);
  uint32_t* ss = reinterpret_cast<uint32_t*>(h->h_live + off_seed);
  *reinterpret_cast<double*>(h->h_live + off_time) = time_to_predict;
  host_parallel_for(h, T, [&](int t) {
    const int n = counts[t * cstride];
    std::memcpy(sb + (size_t)t * D * 2, image_points + (size_t)t * stride, sizeof(double) * 2 * (size_t)n);
    sc[t] = n; ss[t] = seeds[t * cstride];
  });
  static cudaEvent_t eA = nullptr, eB = nullptr;
  if (trace && !eA) { cudaEventCreate(&eA); cudaEventCreate(&eB); }
  if (trace) { t1 = now(); cudaEventRecord(eA, h->stream); }
  PF_CUDA(h, cudaMemcpyAsync(h->live.blobs, sb, nb, cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(h->live.times, h->h_live + off_time, 8, cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(h->live.counts, sc, (size_t)T * 4, cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(h->live.seeds, ss, (size_t)T * 4, cudaMemcpyHostToDevice, h->stream));
  std::swap(h->seq, h->live);                      // the track-mode launch reads its sequence from h->seq
  {
    pf_kparams p;
    fill_kparams(h, p);
    p.blob_count = h->seq.counts; p.times = h->seq.times; p.seeds = h->seq.seeds; p.total_frames = 1;
    pf_track_prepare_kernel<<<(T + 127) / 128, 128, 0, h->stream>>>(p, 0);      // predictPose (:234 -> :995-1010) per tracker
    rc = cudaGetLastError() == cudaSuccess ? PFMPE_OK : PFMPE_ERR_CUDA;
    h->launches += 1;
    // Gauss-Newton on the device for at most gn_cap iterations (PFMPE_GN_CAP, default 64: about one object in 110 goes beyond on the bench scene; 0 = the full 500): objects
    // still iterating then are finished by the host below
    static const int cap_env = [] { const char* e = std::getenv("PFMPE_GN_CAP"); return e ? std::atoi(e) : 64; }();
    h->gn_cap = cap_env;
    if (rc == PFMPE_OK) rc = launch_frames(h, 1, 0, T, 0, 1);
    h->gn_cap = 0;
  }
  std::swap(h->seq, h->live);
  if (rc) return rc;
  // the capped objects first, packed, straight into pinned memory; then the full record array
  pf_track_late_kernel<<<1, 1024, 0, h->stream>>>(h->live.result, T, h->h_late_hdr, h->h_late_rec);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  PF_CUDA(h, cudaEventRecord(h->ev_late, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(h->h_live_res, h->live.result, (size_t)T * sizeof(pf_frame_result), cudaMemcpyDeviceToHost, h->stream));
  if (trace) { cudaEventRecord(eB, h->stream); t2 = now(); }
  PF_CUDA(h, cudaEventSynchronize(h->ev_late));
  if (trace) t3 = now();
  // objects whose device Gauss-Newton was capped: the whole optimisation again on the host (:690 -> :1805-2009, full
  // semantics), in parallel over the objects and under the copy of the full array; their records and device state
  // (cur / pred / cov) are patched
  auto redo = [&](pf_frame_result& s, int t) {
    pf_pose Tw;
    for (int i = 0; i < 12; ++i) Tw.p[i] = s.winner_pose[i];
    pf_aff pose = pf_aff_from_f32(Tw);
    const pf_aff before = pose;
    double c[36];
    s.gn_iters = pf_optimise_pose(h->K, &h->markers[(size_t)t * PF_MAX_LED * 3], image_points + (size_t)t * stride, s.pairs, s.npairs, pose, c);
    s.flag_fail = 1.0f;
    for (int r = 0; r < 3; ++r)
      for (int k = 0; k < 3; ++k)
        if (fabs(before.a[4 * r + k] - pose.a[4 * r + k]) >= 0.3) s.flag_fail = 1.5f;   // :693-701
    for (int i = 0; i < 12; ++i) s.pose[i] = pose.a[i];
    for (int i = 0; i < 36; ++i) s.cov[i] = c[i];
  };
  auto redo_all = [&](int n, const std::function<void(int)>& one) {
    if (n >= 2 && h->pool && h->pool->size() > 0) h->pool->run_uneven(n, one);
    else for (int q = 0; q < n; ++q) one(q);
  };
  int* list = h->h_late_hdr + 4;
  int n_late = h->h_late_hdr[0];
  if (n_late > 0) redo_all(n_late, [&](int q) { redo(h->h_late_rec[q], list[q]); });
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  float pre_ms = 0.f, post_ms = 0.f;
  if (trace) { cudaEventElapsedTime(&pre_ms, eA, h->ev0); cudaEventElapsedTime(&post_ms, h->ev1, eB); }
  if (n_late > 0) {
    for (int q = 0; q < n_late; ++q) h->h_live_res[list[q]] = h->h_late_rec[q];
  } else if (n_late < 0) {                          // more than PF_LATE_MAX of them: out of the full array
    n_late = 0;
    for (int t = 0; t < T; ++t)
      if (h->h_live_res[t].status != 0 && h->h_live_res[t].resampled && h->h_live_res[t].gn_iters == -2) list[n_late++] = t;
    redo_all(n_late, [&](int q) { redo(h->h_live_res[list[q]], list[q]); });
  }
  if (n_late > 0) {
    pf_track_patch_kernel<<<(unsigned)n_late, 64, 0, h->stream>>>(h->state, h->h_live_res, list, n_late);
    PF_CUDA(h, cudaGetLastError());
    h->launches += 1;
    PF_CUDA(h, cudaStreamSynchronize(h->stream));   // the pinned records and the list are rewritten by the next frame
  }
  if (trace) t4 = now();
  host_parallel_for(h, T, [&](int t) {
    const pf_frame_result& s = h->h_live_res[t];
    pfmpe_track_result* o = out + (size_t)t * ostride;
    std::memset(o, 0, sizeof *o);
    o->status = s.status; o->flag_fail = s.flag_fail; o->iterations = s.iterations; o->gn_iterations = s.gn_iters;
    o->most_likely_pose_idx = s.most_likely_pose_idx; o->max_count = s.max_count;
    o->highest_prob = s.highest_prob; o->weight_sum = (double)s.sum_fx / 1073741824.0;
    pf_aff P;
    for (int i = 0; i < 12; ++i) P.a[i] = s.pose[i];
    pf_aff_to_cm16(P, o->pose);
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) o->cov[i + 6 * j] = s.cov[6 * i + j];
    pf_aff W;
    for (int i = 0; i < 12; ++i) W.a[i] = (double)s.winner_pose[i];
    pf_aff_to_cm16(W, o->winner_pose);
    o->n_correspondences = s.npairs;
    for (int i = 0; i < 2 * PFMPE_MAX_LED; ++i) o->correspondences[i] = s.pairs[i];
    h->last_sum_fx[t] = s.sum_fx;
    // host mirror of the tracked object (updatePose :2011-2021), so that the host-tail calls can take over
    HostTrack& ht = h->host_trk[t];
    if (s.status != 0) {
      if (s.resampled) {
        if (ht.isi < 2) ++ht.isi;
        ht.prev = ht.cur; ht.cur = P;
        if (time_to_predict - ht.t_cur > 0.001 || time_to_predict < ht.t_cur) { ht.t_prev = ht.t_cur; ht.t_cur = time_to_predict; }
        for (int i = 0; i < 36; ++i) ht.cov[i] = s.cov[i];
      } else ht.isi = 0;
    }
  });
  h->last_was_track = false;
  if (trace) {
    const double t5 = now();
    float kms = 0.f;
    cudaEventElapsedTime(&kms, h->ev0, h->ev1);
    tr.push_back({t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, (double)kms * 1e-3, (double)n_late, t5 - t0, (double)pre_ms * 1e-3, (double)post_ms * 1e-3});
    if (tr.size() % 20 == 0) {
      const char* nm[10] = {"stage", "enqueue", "wait (capped list)", "host GN + rest of D2H", "convert", "kernel(events)", "late objects", "total", "H2D+prepare(events)", "D2H(events)"};
      std::fprintf(stderr, "[pfmpe trace, last 20 frames, median / max]");
      for (int k = 0; k < 10; ++k) {
        std::vector<double> v;
        for (size_t i = tr.size() - 20; i < tr.size(); ++i) v.push_back(tr[i][(size_t)k]);
        std::sort(v.begin(), v.end());
        if (k == 6) std::fprintf(stderr, "  %s %.0f / %.0f", nm[k], v[10], v[19]);
        else std::fprintf(stderr, "  %s %.3f / %.3f ms", nm[k], 1e3 * v[10], 1e3 * v[19]);
      }
      std::fprintf(stderr, "\n");
    }
  }
  return PFMPE_OK;
}

int pfmpe_track_frame_device_batch(pfmpe_handle* h, double time_to_predict, const double* image_points, const int32_t* counts,
                                   const uint32_t* seeds, pfmpe_track_result* out)
{
  if (!h || !out || !image_points || !counts || !seeds) return PFMPE_ERR_INVALID;
  return track_frames_device(h, time_to_predict, image_points, (size_t)h->cfg.max_blobs * 2, counts, seeds, 1, out, 1);
}

int pfmpe_track_sequence_device(pfmpe_handle* h, int n_frames, const double* times, const double* image_points,
                                const int32_t* counts, const uint32_t* seeds, pfmpe_track_result* out, double* frame_seconds)
{
  if (!h || n_frames < 1 || !times || !image_points || !counts || !seeds || !out) return PFMPE_ERR_INVALID;
  const size_t D2 = (size_t)h->cfg.max_blobs * 2;
  for (int f = 0; f < n_frames; ++f) {
    const auto a = std::chrono::steady_clock::now();
    const int rc = track_frames_device(h, times[f], image_points + (size_t)f * D2, (size_t)n_frames * D2, counts + f, seeds + f,
                                       (size_t)n_frames, out + f, (size_t)n_frames);
    if (rc) return rc;
    if (frame_seconds) frame_seconds[f] = std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count();
  }
  return PFMPE_OK;
}

// ---- ROI prediction (K8) -------------------------------------------------------
namespace {
typedef void (*roi_kernel_t)(const pf_kparams, const pf_pose, const pf_pose, const pf_pose, unsigned*);
roi_kernel_t roi_kernel_for(int L)
{
  switch (L) {
    case 1: return pf_roi_kernel<1>; case 2: return pf_roi_kernel<2>; case 3: return pf_roi_kernel<3>;
    case 4: return pf_roi_kernel<4>; case 5: return pf_roi_kernel<5>; case 6: return pf_roi_kernel<6>;
    case 7: return pf_roi_kernel<7>; case 8: return pf_roi_kernel<8>; default: return nullptr;
  }
}
}  // namespace

int pfmpe_predict_roi(pfmpe_handle* h, int tracker, const double cam_move_inv[16], const double prediction_matrix[16],
                      const double predicted_pose[16], double roi[4])
{
  if (!h || !cam_move_inv || !prediction_matrix || !predicted_pose || !roi) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  int rc = check_ready(h);
  if (rc) return rc;
  if (!h->particles_set[tracker]) return fail(h, PFMPE_ERR_STATE, "set_particles / track_init must precede predict_roi");
  roi_kernel_t k = roi_kernel_for(h->cfg.n_leds);
  if (!k) return fail(h, PFMPE_ERR_UNSUPPORTED, "n_leds must be 1..8");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  pf_kparams p;
  fill_kparams(h, p);
  p.tracker_base = tracker;
  const pf_pose Cm = pf_aff_to_f32(pf_aff_from_cm16(cam_move_inv)), Pm = pf_aff_to_f32(pf_aff_from_cm16(prediction_matrix)),
                pr = pf_aff_to_f32(pf_aff_from_cm16(predicted_pose));
  const int N = h->cfg.n_particles;
  int blocks = (N + PF_BLOCK) / PF_BLOCK;
  if (blocks > h->num_sms * 8) blocks = h->num_sms * 8;
  unsigned* keys = h->roi_keys + 4 * (size_t)tracker;
  pf_roi_init_kernel<<<1, 32, 0, h->stream>>>(keys, 1);
  k<<<dim3((unsigned)blocks, 1, 1), PF_BLOCK, 0, h->stream>>>(p, Cm, Pm, pr, h->roi_keys);
  pf_roi_finish_kernel<<<1, 32, 0, h->stream>>>(keys, h->roi_out + 4 * (size_t)tracker, 4);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 3;
  PF_CUDA(h, cudaMemcpyAsync(h->h_roi + 4 * (size_t)tracker, h->roi_out + 4 * (size_t)tracker, 4 * sizeof(float),
                             cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  for (int i = 0; i < 4; ++i) roi[i] = (double)h->h_roi[4 * (size_t)tracker + i];
  return PFMPE_OK;
}

// ---- K9: correspondence vote of initialise() ----------------------------------
namespace {
int p3p_histogram_impl(pfmpe_handle* h, int tracker, const double* image_points, bool on_device, int32_t D,
                       uint32_t* hist, pfmpe_p3p_stats* stats)
{
  if (!h || !image_points || !hist) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  int rc = check_ready(h);
  if (rc) return rc;
  const int L = h->cfg.n_leds;
  if (L < 3) return fail(h, PFMPE_ERR_UNSUPPORTED, "P3P needs at least 3 LEDs");
  if (D < 3 || D > PF_P3P_MAX_D) return fail(h, PFMPE_ERR_INVALID, "n_image_points must be 3..1024");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  // scratch layout: img 2D | simg 2D | vec 3D | stats 4 | wtab 2*nperm*16 (all 8-byte) | hist D*L u32 | sidx D i32
  const int nperm = L * (L - 1) * (L - 2);
  const size_t wtab_doubles = (size_t)2 * nperm * (sizeof(pf_p3p_world) / sizeof(double));
  const size_t need = ((size_t)D * 7 + 4 + wtab_doubles) * 8 + (size_t)D * L * 4 + (size_t)D * 4;
  if (need > h->p3p_bytes) {
    PF_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaFree(h->p3p_buf); h->p3p_buf = nullptr; h->p3p_bytes = 0;
    PF_CUDA(h, cudaMalloc((void**)&h->p3p_buf, need));
    h->p3p_bytes = need;
  }
  pf_p3p_params p;
  std::memset(&p, 0, sizeof p);
  p.D = D; p.L = L;
  for (int i = 0; i < 9; ++i) p.K[i] = h->K[i];
  for (int i = 0; i < 3 * L; ++i) p.obj[i] = h->markers[(size_t)tracker * PF_MAX_LED * 3 + i];
  p.tol = h->prm.back_projection_pixel_tolerance;
  double* base = reinterpret_cast<double*>(h->p3p_buf);
  double* d_img = base;
  p.simg = base + 2 * (size_t)D;
  p.vec = base + 4 * (size_t)D;
  p.stats = reinterpret_cast<unsigned long long*>(base + 7 * (size_t)D);
  p.wtab = reinterpret_cast<pf_p3p_world*>(base + 7 * (size_t)D + 4);
  p.hist = reinterpret_cast<unsigned*>(base + 7 * (size_t)D + 4 + wtab_doubles);
  // enough threads to fill the GPU even for a handful of detections: slice the permutation list of a triple
  const long long triples = (long long)D * (D - 1) * (D - 2) / 6;
  const long long want_threads = (long long)h->num_sms * 3 * PF_P3P_BLOCK * 2;
  long long nchunk = (want_threads + triples - 1) / triples;
  p.nchunk = (int)(nchunk < 1 ? 1 : nchunk > nperm ? nperm : nchunk);
  p.sidx = reinterpret_cast<int*>(p.hist + (size_t)D * L);
  if (on_device) p.img = image_points;
  else {
    PF_CUDA(h, cudaMemcpyAsync(d_img, image_points, (size_t)D * 2 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    p.img = d_img;
  }
  // Large votes run as two kernels per slab of threads (roots, then back-projection + votes; see pf_p3p_hist_kernel) through a
  // roots buffer of 32 B per P3P problem, at most PFMPE_P3P_ROOTS_MB (default 3072; 0 = always the single kernel) of it, kept
  // with the handle.  Below 2e6 problems the single kernel is as fast and needs no buffer.
  static const long long roots_cap = [] { const char* e = std::getenv("PFMPE_P3P_ROOTS_MB"); return (e ? std::atoll(e) : 3072ll) << 20; }();
  const size_t per_thread = pf_p3p_roots_bytes_per_thread(L, p.nchunk);
  const long long n_threads = triples * p.nchunk;
  size_t want_roots = 0;
  if (roots_cap > 0 && n_threads * (long long)(per_thread / 32) >= 2000000ll) {
    const unsigned long long all = (unsigned long long)n_threads * per_thread;
    want_roots = (size_t)(all < (unsigned long long)roots_cap ? all : (unsigned long long)roots_cap);
  }
  if (want_roots > h->p3p_roots_bytes) {
    PF_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaFree(h->p3p_roots); h->p3p_roots = nullptr; h->p3p_roots_bytes = 0;
    PF_CUDA(h, cudaMalloc((void**)&h->p3p_roots, want_roots));
    h->p3p_roots_bytes = want_roots;
  }
  PF_CUDA(h, cudaEventRecord(h->ev0, h->stream));
  PF_CUDA(h, pf_p3p_launch_hist(p, h->stream, want_roots ? h->p3p_roots : nullptr, want_roots ? h->p3p_roots_bytes : 0));
  PF_CUDA(h, cudaEventRecord(h->ev1, h->stream));
  h->launches += 2;
  h->timing_valid = false;
  unsigned long long st[4];
  PF_CUDA(h, cudaMemcpyAsync(hist, p.hist, (size_t)D * L * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(st, p.stats, sizeof st, cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  if (stats) {
    stats->triples = st[0]; stats->p3p_solved = st[1]; stats->solutions_projected = st[2]; stats->solutions_voted = st[3];
    stats->reserved = 0;
    PF_CUDA(h, cudaEventElapsedTime(&stats->device_ms, h->ev0, h->ev1));
  }
  return PFMPE_OK;
}
}  // namespace

int pfmpe_p3p_histogram(pfmpe_handle* h, int tracker, const double* image_points, int32_t n_image_points,
                        uint32_t* hist, pfmpe_p3p_stats* stats)
{
  return p3p_histogram_impl(h, tracker, image_points, false, n_image_points, hist, stats);
}

int pfmpe_p3p_histogram_dev(pfmpe_handle* h, int tracker, const double* d_image_points, int32_t n_image_points,
                            uint32_t* hist, pfmpe_p3p_stats* stats)
{
  return p3p_histogram_impl(h, tracker, d_image_points, true, n_image_points, hist, stats);
}

// ---- the vote without any allocation: scratch in a file-scope device array --------------------------------------
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels as pfmpe_p3p_histogram
// (pf_p3p_prep_kernel + pf_p3p_hist_kernel); their scratch -- the device copy of the detections, the copy sorted by x,
// the bearing vectors, the world-frame table of the LED permutations, the counters and the histogram -- lives in the
// static array below instead of the handle's grow-on-demand cudaMalloc buffer.  Every GPU that loads the library holds
// the array until the process exits.  All calls on a device share it and it keeps the last call's contents: the
// prologue kernel rewrites every table and zeroes the histogram and the counters at the start of each call, so calls
// that are ordered on the GPU are independent; calls that OVERLAP on the GPU (two handles / two streams at once) race.
#define PF_STATIC_SCRATCH_BYTES (128 * 1024)
static __device__ __align__(256) unsigned char pf_static_scratch[PF_STATIC_SCRATCH_BYTES];

namespace {
// In host code `pf_static_scratch` is a host-side stand-in, not a device address: the address is asked for once per
// device, with that device current, and cached.
std::mutex g_static_mu;
unsigned char* g_static_ptr[64] = {nullptr};

int static_scratch_ptr(pfmpe_handle* h, unsigned char** out)
{
  const int dev = h->cfg.device;
  if (dev < 0 || dev >= 64) return fail(h, PFMPE_ERR_UNSUPPORTED, "device ordinal beyond the static scratch table (64)");
  std::lock_guard<std::mutex> g(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* q = nullptr;
    PF_CUDA(h, cudaSetDevice(dev));
    PF_CUDA(h, cudaGetSymbolAddress(&q, pf_static_scratch));
    g_static_ptr[dev] = static_cast<unsigned char*>(q);
  }
  *out = g_static_ptr[dev];
  return PFMPE_OK;
}
}  // namespace

int pfmpe_p3p_histogram_static(pfmpe_handle* h, int tracker, const double* image_points, int32_t D, uint32_t* hist,
                               pfmpe_p3p_stats* stats)
{
  if (!h || !image_points || !hist) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  int rc = check_ready(h);
  if (rc) return rc;
  const int L = h->cfg.n_leds;
  if (L < 3) return fail(h, PFMPE_ERR_UNSUPPORTED, "P3P needs at least 3 LEDs");
  if (D < 3) return fail(h, PFMPE_ERR_INVALID, "n_image_points must be >= 3");
  if (D > PFMPE_P3P_STATIC_MAX_POINTS)
    return fail(h, PFMPE_ERR_INVALID, "pfmpe_p3p_histogram_static: n_image_points exceeds PFMPE_P3P_STATIC_MAX_POINTS (256); "
                                      "use pfmpe_p3p_histogram for larger frames");
  // scratch layout (as pfmpe_p3p_histogram): img 2D | simg 2D | vec 3D | stats 4 | wtab 2*nperm*16 (all 8-byte) | hist D*L u32 | sidx D i32
  const int nperm = L * (L - 1) * (L - 2);
  const size_t wtab_doubles = (size_t)2 * nperm * (sizeof(pf_p3p_world) / sizeof(double));
  const size_t need = ((size_t)D * 7 + 4 + wtab_doubles) * 8 + (size_t)D * L * 4 + (size_t)D * 4;
  if (need > PF_STATIC_SCRATCH_BYTES) return fail(h, PFMPE_ERR_INVALID, "pfmpe_p3p_histogram_static: problem exceeds the static scratch (128 KiB)");
  unsigned char* scratch = nullptr;
  rc = static_scratch_ptr(h, &scratch);
  if (rc) return rc;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  pf_p3p_params p;
  std::memset(&p, 0, sizeof p);
  p.D = D; p.L = L;
  for (int i = 0; i < 9; ++i) p.K[i] = h->K[i];
  for (int i = 0; i < 3 * L; ++i) p.obj[i] = h->markers[(size_t)tracker * PF_MAX_LED * 3 + i];
  p.tol = h->prm.back_projection_pixel_tolerance;
  double* base = reinterpret_cast<double*>(scratch);
  double* d_img = base;
  p.img = d_img;
  p.simg = base + 2 * (size_t)D;
  p.vec = base + 4 * (size_t)D;
  p.stats = reinterpret_cast<unsigned long long*>(base + 7 * (size_t)D);
  p.wtab = reinterpret_cast<pf_p3p_world*>(base + 7 * (size_t)D + 4);
  p.hist = reinterpret_cast<unsigned*>(base + 7 * (size_t)D + 4 + wtab_doubles);
  p.sidx = reinterpret_cast<int*>(p.hist + (size_t)D * L);
  const long long triples = (long long)D * (D - 1) * (D - 2) / 6;
  const long long want_threads = (long long)h->num_sms * 3 * PF_P3P_BLOCK * 2;
  long long nchunk = (want_threads + triples - 1) / triples;
  p.nchunk = (int)(nchunk < 1 ? 1 : nchunk > nperm ? nperm : nchunk);
  PF_CUDA(h, cudaMemcpyAsync(d_img, image_points, (size_t)D * 2 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaEventRecord(h->ev0, h->stream));
  PF_CUDA(h, pf_p3p_launch_hist(p, h->stream));   // the prologue zeroes hist / stats and rewrites every table
  PF_CUDA(h, cudaEventRecord(h->ev1, h->stream));
  h->launches += 2;
  h->timing_valid = false;
  unsigned long long st[4];
  PF_CUDA(h, cudaMemcpyAsync(hist, p.hist, (size_t)D * L * sizeof(uint32_t), cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(st, p.stats, sizeof st, cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  if (stats) {
    stats->triples = st[0]; stats->p3p_solved = st[1]; stats->solutions_projected = st[2]; stats->solutions_voted = st[3];
    stats->reserved = 0;
    PF_CUDA(h, cudaEventElapsedTime(&stats->device_ms, h->ev0, h->ev1));
  }
  return PFMPE_OK;
}

int pfmpe_p3p_check(pfmpe_handle* h, int tracker, const double* image_points, int32_t D, const uint32_t* correspondences,
                    const int32_t* counts, int32_t M, int32_t stride, double certainty_threshold, pfmpe_p3p_check_result* results)
{
  static_assert(sizeof(pfmpe_p3p_check_result) == sizeof(pf_p3p_check_rec) && PFMPE_P3P_CHECK_STRIDE == PF_P3P_CHECK_STRIDE,
                "ABI record and kernel record must agree");
  if (!h || !image_points || !correspondences || !counts || !results) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  int rc = check_ready(h);
  if (rc) return rc;
  const int L = h->cfg.n_leds;
  if (L < 3) return fail(h, PFMPE_ERR_UNSUPPORTED, "P3P needs at least 3 LEDs");
  if (D < 3 || D > PF_P3P_MAX_D || M < 0 || M > (1 << 22)) return fail(h, PFMPE_ERR_INVALID, "n_image_points must be 3..1024, n_sets 0..2^22");
  if (M == 0) return PFMPE_OK;
  if (stride < 1 || stride > PF_P3P_CHECK_STRIDE) return fail(h, PFMPE_ERR_INVALID, "stride must be 1..56");
  for (int s = 0; s < M; ++s) {
    const int n = counts[s];
    if (n >= 3 && n <= PF_P3P_MAX_L && n * (n - 1) * (n - 2) / 6 > stride) return fail(h, PFMPE_ERR_INVALID, "stride < C(counts[s], 3)");
  }
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  // scratch: img 2D doubles | results M*stride records | corr M*16 u32 | counts M i32
  const size_t rec_bytes = (size_t)M * stride * sizeof(pf_p3p_check_rec);
  const size_t need = (size_t)D * 16 + rec_bytes + (size_t)M * (2 * PF_P3P_MAX_L * 4 + 4);
  if (need > h->p3p_io_bytes) {
    PF_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaFree(h->p3p_io); h->p3p_io = nullptr; h->p3p_io_bytes = 0;
    PF_CUDA(h, cudaMalloc((void**)&h->p3p_io, need));
    h->p3p_io_bytes = need;
  }
  double* d_img = h->p3p_io;
  pf_p3p_check_rec* d_rec = reinterpret_cast<pf_p3p_check_rec*>(d_img + 2 * (size_t)D);
  unsigned* d_corr = reinterpret_cast<unsigned*>(reinterpret_cast<unsigned char*>(d_rec) + rec_bytes);
  int* d_cnt = reinterpret_cast<int*>(d_corr + (size_t)M * 2 * PF_P3P_MAX_L);
  pf_p3p_params p;
  std::memset(&p, 0, sizeof p);
  p.D = D; p.L = L;
  for (int i = 0; i < 9; ++i) p.K[i] = h->K[i];
  for (int i = 0; i < 3 * L; ++i) p.obj[i] = h->markers[(size_t)tracker * PF_MAX_LED * 3 + i];
  p.tol = h->prm.back_projection_pixel_tolerance;
  p.img = d_img;
  PF_CUDA(h, cudaMemcpyAsync(d_img, image_points, (size_t)D * 16, cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(d_corr, correspondences, (size_t)M * 2 * PF_P3P_MAX_L * 4, cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(d_cnt, counts, (size_t)M * 4, cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaMemsetAsync(d_rec, 0, rec_bytes, h->stream));
  PF_CUDA(h, pf_p3p_launch_check(p, M, stride, d_corr, d_cnt, certainty_threshold, d_rec, h->stream));
  h->launches += 1;
  PF_CUDA(h, cudaMemcpyAsync(results, d_rec, rec_bytes, cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  return PFMPE_OK;
}

int pfmpe_p3p_solve(pfmpe_handle* h, int32_t n, const double* feature_vectors, const double* world_points,
                    double* solutions, int32_t* rc_out)
{
  if (!h || !feature_vectors || !world_points || !solutions || !rc_out) return PFMPE_ERR_INVALID;
  if (n < 0) return fail(h, PFMPE_ERR_INVALID, "n < 0");
  if (n == 0) return PFMPE_OK;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  const size_t need = (size_t)n * (9 + 9 + 48) * sizeof(double) + (size_t)n * sizeof(int);
  if (need > h->p3p_io_bytes) {
    PF_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaFree(h->p3p_io); h->p3p_io = nullptr; h->p3p_io_bytes = 0;
    PF_CUDA(h, cudaMalloc((void**)&h->p3p_io, need));
    h->p3p_io_bytes = need;
  }
  double* d_fv = h->p3p_io;
  double* d_wp = d_fv + 9 * (size_t)n;
  double* d_sol = d_wp + 9 * (size_t)n;
  int* d_rc = reinterpret_cast<int*>(d_sol + 48 * (size_t)n);
  PF_CUDA(h, cudaMemcpyAsync(d_fv, feature_vectors, (size_t)n * 9 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(d_wp, world_points, (size_t)n * 9 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  PF_CUDA(h, cudaMemsetAsync(d_sol, 0, (size_t)n * 48 * sizeof(double), h->stream));
  PF_CUDA(h, pf_p3p_launch_solve(n, d_fv, d_wp, d_sol, d_rc, h->stream));
  h->launches += 1;
  PF_CUDA(h, cudaMemcpyAsync(solutions, d_sol, (size_t)n * 48 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaMemcpyAsync(rc_out, d_rc, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  return PFMPE_OK;
}

// ---- sharded single filter (BASELINE configs[4]; one handle per rank) ---------------------------
// The particle set of ONE filter is split into contiguous shards, one per handle/GPU.  Every stage
// is one launch; the exchange steps between them (all-gather of the 64-byte partials, all-to-all
// of the resampled rows) belong to the caller (pf_monocular_pose_estimator_b200/sharded.py: NCCL).
int pfmpe_shard_configure(pfmpe_handle* h, int64_t global_offset, int64_t n_global)
{
  if (!h) return PFMPE_ERR_INVALID;
  if (h->cfg.n_trackers != 1) return fail(h, PFMPE_ERR_UNSUPPORTED, "a sharded filter is one tracked object per handle");
  const int64_t N = h->cfg.n_particles;
  if (global_offset < 0 || n_global < 2 || global_offset + N > n_global || n_global > 0x7fffffff)
    return fail(h, PFMPE_ERR_INVALID, "shard range outside [0, n_global) or n_global >= 2^31");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  // a shard's sources can own any number of the n_global output slots
  cudaFree(h->idx); h->idx = nullptr;
  PF_CUDA(h, dev_alloc(&h->idx, (size_t)n_global));
  PF_CUDA(h, cudaMemset(h->idx, 0, (size_t)n_global * sizeof(int)));
  // One allocation [pose0: 3 planes x 2N | pose1: 3 planes x 2N | lead: N] -- one CUDA IPC handle for the peers.  The
  // second half of every plane is the receive area the other ranks' rebalance writes rows into.
  if (h->particles_set[0]) return fail(h, PFMPE_ERR_STATE, "pfmpe_shard_configure must precede set_particles / track_init");
  cudaFree(h->pose0); if (!h->shard_alloc) cudaFree(h->pose1);
  h->pose0 = h->pose1 = nullptr; h->lead = nullptr;
  const size_t S = 2 * (size_t)N, bytes = 6 * S * sizeof(float4) + (size_t)N * sizeof(int);
  PF_CUDA(h, cudaMalloc((void**)&h->pose0, bytes));
  PF_CUDA(h, cudaMemset(h->pose0, 0, bytes));
  h->shard_alloc = true;
  h->pstride = S;
  h->pose1 = h->pose0 + 3 * S;
  h->lead = reinterpret_cast<int*>(h->pose0 + 6 * S);
  h->sh_src = 0;
  if (!h->warp_agg_g) PF_CUDA(h, dev_alloc(&h->warp_agg_g, (size_t)PF_MAX_GRID * PF_MAX_WARPS));
  if (!h->shard_partial) PF_CUDA(h, dev_alloc(&h->shard_partial, PF_XMAIL_WORDS));
  h->sharded = true;
  h->goff = (int)global_offset;
  h->n_global = (int)n_global;
  return PFMPE_OK;
}

static int shard_ready(pfmpe_handle* h)
{
  if (!h->sharded) return fail(h, PFMPE_ERR_STATE, "pfmpe_shard_configure must be called first");
  int rc = check_ready(h);
  if (rc) return rc;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  return PFMPE_OK;
}

int pfmpe_shard_begin_frame(pfmpe_handle* h, const pfmpe_frame_in* in)
{
  if (!h || !in) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  PF_CUDA(h, cudaStreamSynchronize(h->stream));   // the pinned inbox may still be in flight
  rc = check_frame(h, 0, in);
  if (rc) return rc;
  stage_frame(h, 0, in);
  PF_CUDA(h, cudaMemcpyAsync(h->inbox, h->h_inbox, h->inbox_bytes, cudaMemcpyHostToDevice, h->stream));
  return PFMPE_OK;
}

int pfmpe_shard_score(pfmpe_handle* h, int iteration, void* d_partial)
{
  if (!h || iteration < 0 || iteration >= PF_MAX_ITER) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  ShardArgs sa;
  sa.pass_it = iteration;
  sa.partial = static_cast<long long*>(d_partial);
  return launch_frames(h, 3, 0, 1, 0, 1, &sa);
}

int pfmpe_shard_resample(pfmpe_handle* h, int64_t sum_fx, int iterations, const int64_t lower[3], const int64_t all[3],
                         void* d_partial)
{
  if (!h || !lower || !all || iterations < 1 || iterations > PF_MAX_ITER) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  ShardArgs sa;
  sa.S_fx = sum_fx;
  sa.iters_total = iterations;
  for (int i = 0; i < 3; ++i) { sa.excl[i] = lower[i]; sa.total[i] = all[i]; }
  sa.partial = static_cast<long long*>(d_partial);
  rc = launch_frames(h, 4, 0, 1, 0, 1, &sa);
  if (rc == PFMPE_OK && h->peers_dev) h->sh_src ^= 1;   // peer-memory rebalance: the propagated set becomes the source (see the kernel's mode-4 epilogue)
  return rc;
}

int pfmpe_shard_export(pfmpe_handle* h, int64_t first_owned_slot, int64_t first_slot, int64_t count, void* d_rows)
{
  if (!h || count < 0 || (count > 0 && !d_rows) || first_slot < first_owned_slot) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  if (count == 0) return PFMPE_OK;
  const size_t N = (size_t)h->cfg.n_particles;
  if (first_slot - first_owned_slot + count > (int64_t)h->n_global)
    return fail(h, PFMPE_ERR_INVALID, "slot range outside this shard's output");
  const int blocks = (int)((count + PF_BLOCK - 1) / PF_BLOCK);
  pf_shard_export_kernel<<<blocks, PF_BLOCK, 0, h->stream>>>(h->sh_src == 0 ? h->pose1 : h->pose0, h->idx, (int)h->pstride, first_owned_slot, first_slot, count,
                                                            1, static_cast<float4*>(d_rows));
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  return PFMPE_OK;
}

int pfmpe_shard_import(pfmpe_handle* h, const void* d_rows, int64_t local_offset, int64_t count)
{
  if (!h || count < 0 || (count > 0 && !d_rows)) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  const size_t N = (size_t)h->cfg.n_particles;
  if (local_offset < 0 || local_offset + count > (int64_t)N) return fail(h, PFMPE_ERR_INVALID, "rows outside this shard");
  h->particles_set[0] = 1;
  if (count == 0) return PFMPE_OK;
  const int blocks = (int)((count + PF_BLOCK - 1) / PF_BLOCK);
  pf_shard_import_kernel<<<blocks, PF_BLOCK, 0, h->stream>>>(static_cast<const float4*>(d_rows), (int)h->pstride, (int)local_offset,
                                                            (int)count, h->sh_src == 0 ? h->pose0 : h->pose1);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  return PFMPE_OK;
}

int pfmpe_shard_keep(pfmpe_handle* h, int64_t first_owned_slot, int64_t first_slot, int64_t count, int64_t local_offset)
{
  if (!h || count < 0 || first_slot < first_owned_slot) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  const size_t N = (size_t)h->cfg.n_particles;
  if (local_offset < 0 || local_offset + count > (int64_t)N) return fail(h, PFMPE_ERR_INVALID, "slots outside this shard");
  if (first_slot - first_owned_slot + count > (int64_t)h->n_global) return fail(h, PFMPE_ERR_INVALID, "slot range outside this shard's output");
  if (count == 0) return PFMPE_OK;
  const int blocks = (int)((count + PF_BLOCK - 1) / PF_BLOCK);
  pf_shard_keep_kernel<<<blocks, PF_BLOCK, 0, h->stream>>>(h->sh_src == 0 ? h->pose1 : h->pose0, h->idx, (int)h->pstride, first_owned_slot, first_slot, (int)count,
                                                          (int)local_offset, h->sh_src == 0 ? h->pose0 : h->pose1);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  return PFMPE_OK;
}

// ---- peer-memory rebalance (NVLink P2P stores instead of the row exchange) -----------------------
int pfmpe_shard_local_ptr(pfmpe_handle* h, void** d_ptr)
{
  if (!h || !d_ptr) return PFMPE_ERR_INVALID;
  *d_ptr = h->pose0;
  return PFMPE_OK;
}

int pfmpe_shard_ipc_export(pfmpe_handle* h, unsigned char handle[64])
{
  if (!h || !handle) return PFMPE_ERR_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  cudaIpcMemHandle_t m;
  PF_CUDA(h, cudaIpcGetMemHandle(&m, h->pose0));
  std::memcpy(handle, &m, 64);
  return PFMPE_OK;
}

int pfmpe_shard_ipc_open(pfmpe_handle* h, const unsigned char handle[64], void** d_ptr)
{
  if (!h || !handle || !d_ptr) return PFMPE_ERR_INVALID;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  cudaIpcMemHandle_t m;
  std::memcpy(&m, handle, 64);
  void* q = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&q, m, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    h->err = std::string("cudaIpcOpenMemHandle: ") + cudaGetErrorString(e);
    return PFMPE_ERR_COMM;
  }
  h->ipc_opened.push_back(q);
  *d_ptr = q;
  return PFMPE_OK;
}

int pfmpe_shard_set_peers(pfmpe_handle* h, int world, void* const* d_ptrs)
{
  if (!h || world < 0 || world > 1024 || (world > 0 && !d_ptrs)) return PFMPE_ERR_INVALID;
  if (!h->sharded) return fail(h, PFMPE_ERR_STATE, "pfmpe_shard_configure must be called first");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  if (world == 0) {   // back to the row exchange (full rows at every slot, no lead[] indirection)
    PF_CUDA(h, cudaStreamSynchronize(h->stream));
    cudaFree(h->peers_dev); h->peers_dev = nullptr; h->peer_world = 0;
    return PFMPE_OK;
  }
  if ((int64_t)world * h->cfg.n_particles != (int64_t)h->n_global) return fail(h, PFMPE_ERR_INVALID, "peer rebalance needs equal shards");
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  cudaFree(h->peers_dev); h->peers_dev = nullptr;
  PF_CUDA(h, cudaMalloc((void**)&h->peers_dev, sizeof(float4*) * (size_t)world));
  PF_CUDA(h, cudaMemcpy(h->peers_dev, d_ptrs, sizeof(float4*) * (size_t)world, cudaMemcpyHostToDevice));
  cudaFree(h->scatter_cursor); h->scatter_cursor = nullptr;
  PF_CUDA(h, cudaMalloc((void**)&h->scatter_cursor, sizeof(unsigned) * (size_t)world));
  h->peer_world = world;
  return PFMPE_OK;
}

int pfmpe_shard_scatter(pfmpe_handle* h, int64_t first_owned_slot, int64_t count)
{
  if (!h || count < 0 || first_owned_slot < 0) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  if (!h->peers_dev) return fail(h, PFMPE_ERR_STATE, "pfmpe_shard_set_peers must be called first");
  if (first_owned_slot + count > (int64_t)h->n_global) return fail(h, PFMPE_ERR_INVALID, "slot range outside the filter");
  h->particles_set[0] = 1;
  if (count == 0) return PFMPE_OK;
  const int N = h->cfg.n_particles;
  long long blocks = (count + PF_BLOCK - 1) / PF_BLOCK;
  const long long cap = (long long)h->num_sms * 32;
  if (blocks > cap) blocks = cap;
  // called after pfmpe_shard_resample: sh_src already names the buffer that holds T (the new source)
  static const bool force_sparse = [] { const char* e = std::getenv("PFMPE_SCATTER"); return e && std::string(e) == "sparse"; }();
  const float4* Tbuf = h->sh_src == 0 ? h->pose0 : h->pose1;
  const int me = (int)(h->goff / N);
  if (N % PF_BLOCK == 0 && h->scatter_cursor && !force_sparse) {   // packed cross-rank rows (pf_shard_scatter_dense_kernel)
    PF_CUDA(h, cudaMemsetAsync(h->scatter_cursor, 0, sizeof(unsigned) * (size_t)h->peer_world, h->stream));
    pf_shard_scatter_dense_kernel<<<(int)blocks, PF_BLOCK, 0, h->stream>>>(Tbuf, h->idx, N, (int)h->pstride, first_owned_slot, count, h->peers_dev,
                                                                          h->sh_src, me, h->scatter_cursor);
  } else {
    pf_shard_scatter_kernel<<<(int)blocks, PF_BLOCK, 0, h->stream>>>(Tbuf, h->idx, h->kend, N, (int)h->pstride, first_owned_slot, count,
                                                                    h->peers_dev, h->sh_src, me);
  }
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  return PFMPE_OK;
}

// ---- peer mailboxes: the 64-byte partials travel by NVLink peer stores from inside the stage kernels ----
int pfmpe_shard_mail_create(pfmpe_handle* h, int world, int rank, void** d_ptr, unsigned char ipc_handle[64])
{
  if (!h || world < 1 || world > 32 || rank < 0 || rank >= world || !d_ptr) return PFMPE_ERR_INVALID;
  if (!h->sharded) return fail(h, PFMPE_ERR_STATE, "pfmpe_shard_configure must be called first");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  if (!h->xmail) {
    const size_t n = (size_t)PF_XMAIL_RING * world * PF_XMAIL_WORDS;
    PF_CUDA(h, dev_alloc(&h->xmail, n));
    PF_CUDA(h, cudaMemset(h->xmail, 0, n * sizeof(long long)));
    const size_t hb = ((size_t)world * PF_XMAIL_WORDS + 1) * sizeof(long long);
    PF_CUDA(h, cudaMallocHost((void**)&h->h_xtable, hb));
    std::memset(h->h_xtable, 0, hb);
  }
  h->xworld = world; h->xrank = rank;
  *d_ptr = h->xmail;
  if (ipc_handle) {
    cudaIpcMemHandle_t m;
    PF_CUDA(h, cudaIpcGetMemHandle(&m, h->xmail));
    std::memcpy(ipc_handle, &m, 64);
  }
  return PFMPE_OK;
}

int pfmpe_shard_mail_set_peers(pfmpe_handle* h, void* const* d_ptrs)
{
  if (!h || !d_ptrs) return PFMPE_ERR_INVALID;
  if (!h->xmail) return fail(h, PFMPE_ERR_STATE, "pfmpe_shard_mail_create must be called first");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  cudaFree(h->xmail_peers_dev); h->xmail_peers_dev = nullptr;
  PF_CUDA(h, cudaMalloc((void**)&h->xmail_peers_dev, sizeof(void*) * (size_t)h->xworld));
  PF_CUDA(h, cudaMemcpy(h->xmail_peers_dev, d_ptrs, sizeof(void*) * (size_t)h->xworld, cudaMemcpyHostToDevice));
  h->xseq = 0;
  return PFMPE_OK;
}

static int mail_ready(pfmpe_handle* h)
{
  if (!h->xmail_peers_dev) return fail(h, PFMPE_ERR_STATE, "pfmpe_shard_mail_set_peers must be called first");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  return PFMPE_OK;
}

int pfmpe_shard_advance(pfmpe_handle* h)
{
  if (!h) return PFMPE_ERR_INVALID;
  int rc = mail_ready(h);
  if (rc) return rc;
  h->xseq += 1;
  return PFMPE_OK;
}

int pfmpe_shard_collect(pfmpe_handle* h, int only_rank, int64_t* table /* [world][PF_XMAIL_WORDS] */)
{
  if (!h || !table) return PFMPE_ERR_INVALID;
  int rc = mail_ready(h);
  if (rc) return rc;
  if (h->xseq == 0 || only_rank >= h->xworld) return fail(h, PFMPE_ERR_STATE, "no exchange is in flight");
  const int W = h->xworld;
  volatile long long* flag = h->h_xtable + (size_t)W * PF_XMAIL_WORDS;
  *flag = 0;
  pf_shard_wait_kernel<<<1, 32, 0, h->stream>>>(h->xmail, (int)(h->xseq % PF_XMAIL_RING), W, only_rank, h->xseq, h->h_xtable);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  for (unsigned spin = 1;; ++spin) {
    const long long v = *flag;
    if (v == h->xseq) break;
    if (v == -1) return fail(h, PFMPE_ERR_COMM, "a peer's message did not arrive (timeout)");
    if ((spin & 0x3fffu) == 0) {
      const cudaError_t q = cudaStreamQuery(h->stream);
      if (q != cudaSuccess && q != cudaErrorNotReady) { h->err = std::string("exchange: ") + cudaGetErrorString(q); return PFMPE_ERR_CUDA; }
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  for (int i = 0; i < W * PF_XMAIL_WORDS; ++i) table[i] = h->h_xtable[i];
  return PFMPE_OK;
}

int pfmpe_shard_signal(pfmpe_handle* h)
{
  if (!h) return PFMPE_ERR_INVALID;
  int rc = mail_ready(h);
  if (rc) return rc;
  h->xseq += 1;
  pf_shard_signal_kernel<<<1, 32, 0, h->stream>>>(h->xmail_peers_dev, h->xworld, h->xrank, (int)(h->xseq % PF_XMAIL_RING), h->xseq);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  return PFMPE_OK;
}

int pfmpe_shard_wait(pfmpe_handle* h)
{
  if (!h) return PFMPE_ERR_INVALID;
  int rc = mail_ready(h);
  if (rc) return rc;
  pf_shard_wait_kernel<<<1, 32, 0, h->stream>>>(h->xmail, (int)(h->xseq % PF_XMAIL_RING), h->xworld, -1, h->xseq, nullptr);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  return PFMPE_OK;
}

int pfmpe_shard_barrier(pfmpe_handle* h)
{
  const int rc = pfmpe_shard_signal(h);
  return rc ? rc : pfmpe_shard_wait(h);
}

int pfmpe_shard_winner_publish(pfmpe_handle* h, int local_index)
{
  if (!h || local_index < 0 || local_index >= h->cfg.n_particles) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  rc = mail_ready(h);
  if (rc) return rc;
  ShardArgs sa;
  sa.q_local = local_index;
  sa.S_fx = 1;
  sa.publish = true;
  return launch_frames(h, 5, 0, 1, 0, 1, &sa);
}

int pfmpe_shard_winner(pfmpe_handle* h, int local_index, pfmpe_frame_out* out)
{
  if (!h || !out || local_index < 0 || local_index >= h->cfg.n_particles) return PFMPE_ERR_INVALID;
  int rc = shard_ready(h);
  if (rc) return rc;
  ShardArgs sa;
  sa.q_local = local_index;
  sa.S_fx = 1;
  rc = launch_frames(h, 5, 0, 1, 0, 1, &sa);
  if (rc) return rc;
  PF_CUDA(h, cudaMemcpyAsync(h->h_result, h->result, sizeof(pf_frame_result), cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  unpack_result(h->h_result[0], out);
  return PFMPE_OK;
}

// ---- the whole sharded update, host logic in C++ ------------------------------------------------------
// Mirrors the reference loop at its global reductions (pose_estimator.cpp:535-616 do/while + max test, :627-633 sum /
// accept, :666-688 resampling + winner): stage launches on every local shard, the 64-byte partials of ALL ranks arrive
// in this process's mailbox by peer stores, and the decisions -- identical on every rank -- are taken here.
namespace {
// K7 on one shard: pass 1 (mean12 == nullptr: sums of w t and w R) or pass 2 (sums of w xi xi^T about the mean pose of the
// whole filter) over this shard's particles, weights normalised by the weight sum S of the whole filter; the K partial sums
// go to every rank's mailbox (one exchange: xseq advances)
int shard_meancov_round(pfmpe_handle* h, const double* mean12, long long S)
{
  int rc = shard_ready(h);
  if (rc) return rc;
  if (!h->mc_part) return fail(h, PFMPE_ERR_STATE, "the handle was created without PFMPE_FLAG_MEAN_COV");
  pf_kparams p;
  fill_kparams(h, p);
  p.tracker_base = 0;
  p.result = h->result;
  const dim3 grid((unsigned)h->mc_blocks, 1, 1);
  const long long S_all = S != 0 ? S : (1ll << 30);    // zero sum: weights are left unnormalised (:628), 2^30 / 2^30 = 1
  int K;
  if (!mean12) {
    pf_meancov_pass1<<<grid, PF_MC_THREADS, 0, h->stream>>>(p, h->mc_part, S_all);
    K = 12;
  } else {
    PF_CUDA(h, cudaMemcpyAsync(h->mc_mean, mean12, 12 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
    pf_meancov_pass2<<<grid, PF_MC_THREADS, 0, h->stream>>>(p, h->mc_mean, h->mc_part, S_all);
    K = 21;
  }
  h->xseq += 1;
  pf_shard_mc_publish_kernel<<<1, 64, 0, h->stream>>>(h->mc_part, h->mc_blocks, K, h->xmail_peers_dev, h->xworld, h->xrank,
                                                      (int)(h->xseq % PF_XMAIL_RING), h->xseq);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 2;
  return PFMPE_OK;
}

inline float shard_unord(unsigned long long key_hi32)
{
  unsigned o = (unsigned)key_hi32;
  unsigned u = (o & 0x80000000u) ? (o ^ 0x80000000u) : ~o;
  float f;
  std::memcpy(&f, &u, 4);
  return f;
}
struct ShardPass {
  unsigned long long maxkey = 0, minkey_inv = 0;
  long long sum = 0;
  std::vector<pf_sm3> aggs;
};
}  // namespace

int pfmpe_shard_frame(pfmpe_handle* const* hs, int n_local, const pfmpe_frame_in* in, pfmpe_frame_out* out)
{
  if (!hs || n_local < 1 || !in || !out) return PFMPE_ERR_INVALID;
  for (int i = 0; i < n_local; ++i) if (!hs[i]) return PFMPE_ERR_INVALID;
  pfmpe_handle* h0 = hs[0];
  const int W = h0->xworld, L = h0->cfg.n_leds;
  const long long Ng = h0->n_global, n_loc = h0->cfg.n_particles;
  for (int i = 0; i < n_local; ++i) {
    pfmpe_handle* h = hs[i];
    if (!h->sharded || !h->xmail_peers_dev || !h->peers_dev)
      return fail(h, PFMPE_ERR_STATE, "pfmpe_shard_frame needs pfmpe_shard_configure, the peer mailboxes and the peer particle planes");
    if (h->xworld != W || h->n_global != Ng || h->cfg.n_particles != n_loc || h->xseq != h0->xseq)
      return fail(h, PFMPE_ERR_STATE, "the local shards do not belong to one filter / are out of step");
    if (n_local > 1 && h->stream != h0->stream)   // stage kernels of co-located shards wait on one another through the mailboxes:
      return fail(h, PFMPE_ERR_STATE, "shards that share a process must share one stream");   // stream order is what makes that safe
  }
  if ((long long)W * n_loc != Ng) return fail(h0, PFMPE_ERR_STATE, "the sharded frame call needs equal shards");
  int rc;
  for (int i = 0; i < n_local; ++i) if ((rc = pfmpe_shard_begin_frame(hs[i], in)) != PFMPE_OK) return rc;
  const int D = in->n_image_points;
  std::vector<int64_t> table((size_t)W * PF_XMAIL_WORDS);

  auto score_pass = [&](int it, ShardPass& r) -> int {
    for (int i = 0; i < n_local; ++i) {
      int e = pfmpe_shard_score(hs[i], it, nullptr);
      if (e) return e;
      hs[i]->shard_passes += 1;
    }
    int e = pfmpe_shard_collect(h0, -1, table.data());
    if (e) return e;
    r = ShardPass();
    r.aggs.resize(W);
    for (int q = 0; q < W; ++q) {
      const int64_t* row = &table[(size_t)q * PF_XMAIL_WORDS];
      const unsigned long long a = (unsigned long long)row[0], b = (unsigned long long)row[1];
      if (a > r.maxkey) r.maxkey = a;
      if (b > r.minkey_inv) r.minkey_inv = b;
      r.sum += row[2];
      r.aggs[q].s = row[3]; r.aggs[q].mx = row[4]; r.aggs[q].mn = row[5];
    }
    return PFMPE_OK;
  };

  // ---- do { propagate, project, score } while (iter < 80 && max < L*min(5,D))  (:535-616)
  const float thr_exit = (float)(L * (D < 5 ? D : 5));
  float highest = 0.0f;
  int best_it = -1, it = 0;
  ShardPass kept, last;
  for (;;) {
    ShardPass r;
    if ((rc = score_pass(it, r)) != PFMPE_OK) return rc;
    const float m = shard_unord(r.maxkey >> 32);
    last = r;
    ++it;
    if (m > highest) { highest = m; best_it = it - 1; kept = r; }       // :608-614 strict >, from 0
    if (it < PF_MAX_ITER && m < thr_exit) continue;                      // :616
    break;
  }
  const int iters = it;
  if (best_it < 0) kept = last;                                          // :619 nothing to restore
  else if (best_it != iters - 1) { if ((rc = score_pass(best_it, kept)) != PFMPE_OK) return rc; }   // regenerate (exact: RNG by address)

  const long long S = kept.sum;
  const bool accept = S != 0 && highest > (float)(L * (D < 3 ? D : 3));  // :633
  const bool resampled = accept && in->it_since_initialized > 0;
  const long long arg_max = (long long)((~kept.maxkey) & 0xFFFFFFFFull), arg_min = (long long)((~kept.minkey_inv) & 0xFFFFFFFFull);
  long long q = S < 0 ? arg_min : arg_max;                               // :714 on the normalised weights
  unsigned max_count = 0;
  if (resampled) {
    std::vector<pf_sm3> lower(W);
    pf_sm3 run = pf_sm3_identity();
    for (int r = 0; r < W; ++r) { lower[r] = run; run = pf_sm3_combine(run, kept.aggs[r]); }
    for (int i = 0; i < n_local; ++i) {
      const int r = hs[i]->xrank;
      const int64_t lo[3] = {lower[r].s, lower[r].mx, lower[r].mn}, all[3] = {run.s, run.mx, run.mn};
      if ((rc = pfmpe_shard_resample(hs[i], S, iters, lo, all, nullptr)) != PFMPE_OK) return rc;
    }
    if ((rc = pfmpe_shard_collect(h0, -1, table.data())) != PFMPE_OK) return rc;
    unsigned long long cntkey = 0;
    std::vector<long long> first(W + 1);
    for (int r = 0; r < W; ++r) {
      const unsigned long long c = (unsigned long long)table[(size_t)r * PF_XMAIL_WORDS];
      if (c > cntkey) cntkey = c;
      first[r] = table[(size_t)r * PF_XMAIL_WORDS + 1];
    }
    first[W] = Ng;      // a rank's slots end where the next rank's begin (its sources own nothing beyond); the last one ends at N
    long long pos = 0;
    for (int r = 0; r < W; ++r) {
      if (first[r + 1] < first[r] || (first[r + 1] > first[r] && first[r] != pos)) return fail(h0, PFMPE_ERR_COMM, "resampled slot ranges do not tile the filter");
      if (first[r + 1] > first[r]) pos = first[r + 1];
    }
    if (pos != Ng) return fail(h0, PFMPE_ERR_COMM, "resampled slot ranges do not cover the filter");
    q = (long long)((~cntkey) & 0xFFFFFFFFull);                          // :686 first max of counterMeas
    max_count = (unsigned)(cntkey >> 32);
    for (int i = 0; i < n_local; ++i) {
      const int r = hs[i]->xrank;
      if ((rc = pfmpe_shard_scatter(hs[i], first[r], first[r + 1] - first[r])) != PFMPE_OK) return rc;
    }
    // every rank's scatter lands before anyone's next score pass: a barrier on the ranks' streams (co-located shards
    // share a stream, which already orders them: all signals first, then one wait)
    for (int i = 0; i < n_local; ++i) if ((rc = pfmpe_shard_signal(hs[i])) != PFMPE_OK) return rc;
    if ((rc = pfmpe_shard_wait(h0)) != PFMPE_OK) return rc;
  }
  // ---- winner pose + correspondences from its owner (:687-691 / :714-716)
  const int owner = (int)(q / n_loc);
  for (int i = 0; i < n_local; ++i) {
    if (hs[i]->xrank == owner) rc = pfmpe_shard_winner_publish(hs[i], (int)(q - (long long)owner * n_loc));
    else rc = pfmpe_shard_advance(hs[i]);
    if (rc) return rc;
  }
  if ((rc = pfmpe_shard_collect(h0, owner, table.data())) != PFMPE_OK) return rc;
  const int64_t* row = &table[(size_t)owner * PF_XMAIL_WORDS];
  std::memset(out, 0, sizeof *out);
  out->status = accept ? 1 : 4;
  out->iterations = iters;
  out->best_iteration = best_it;
  out->resampled = resampled ? 1 : 0;
  out->highest_prob = (double)highest;
  out->weight_sum = (double)S / 1073741824.0;
  out->most_likely_particle_idx = accept ? (best_it >= 0 ? (int)arg_max : -1) : (int)q;
  out->most_likely_pose_idx = resampled ? (int)q : -1;
  out->max_count = max_count;
  {
    float wp[12];
    std::memcpy(wp, row, sizeof wp);
    pf_aff T;
    for (int i = 0; i < 12; ++i) T.a[i] = (double)wp[i];
    pf_aff_to_cm16(T, out->winner_pose);
    uint32_t pr[2 * PF_MAX_LED];
    std::memcpy(pr, row + 6, sizeof pr);
    const int nc = (int)row[6 + PF_MAX_LED];
    out->n_correspondences = nc;
    for (int i = 0; i < 2 * PF_MAX_LED; ++i) out->correspondences[i] = i < 2 * nc ? pr[i] : 0u;
  }
  for (int i = 0; i < n_local; ++i) hs[i]->last_sum_fx[0] = S;
  // ---- K7 (PFMPE_FLAG_MEAN_COV): weighted mean pose + covariance of the whole filter = two exchanges of 12 + 21 doubles
  if (h0->cfg.flags & PFMPE_FLAG_MEAN_COV) {
    double sum12[12], mean[12], sum21[21];
    for (int pass = 0; pass < 2; ++pass) {
      const int K = pass == 0 ? 12 : 21;
      double* acc = pass == 0 ? sum12 : sum21;
      for (int i = 0; i < n_local; ++i) if ((rc = shard_meancov_round(hs[i], pass == 0 ? nullptr : mean, S)) != PFMPE_OK) return rc;
      if ((rc = pfmpe_shard_collect(h0, -1, table.data())) != PFMPE_OK) return rc;
      for (int k = 0; k < K; ++k) acc[k] = 0.0;
      for (int r = 0; r < W; ++r)                       // rank order: the same sum on every rank
        for (int k = 0; k < K; ++k) {
          double v;
          std::memcpy(&v, &table[(size_t)r * PF_XMAIL_WORDS + k], sizeof v);
          acc[k] += v;
        }
      if (pass == 0) {
        pf_mc_closest_rotation(sum12, mean);
        mean[3] = sum12[3]; mean[7] = sum12[7]; mean[11] = sum12[11];
      }
    }
    pf_aff Tm;
    for (int i = 0; i < 12; ++i) Tm.a[i] = mean[i];
    pf_aff_to_cm16(Tm, out->mean_pose);
    int q21 = 0;
    for (int i = 0; i < 6; ++i)
      for (int j = i; j < 6; ++j) { out->mean_cov[6 * i + j] = sum21[q21]; out->mean_cov[6 * j + i] = sum21[q21]; ++q21; }
  }
  return PFMPE_OK;
}

int pfmpe_shard_track_frame_host(pfmpe_handle* const* hs, int n_local, double time_to_predict, const double* image_points,
                                 int32_t n_image_points, uint32_t seed, pfmpe_track_result* out)
{
  if (!hs || n_local < 1 || !hs[0] || !out || (n_image_points > 0 && !image_points)) return PFMPE_ERR_INVALID;
  pfmpe_handle* h = hs[0];
  HostTrack& ht = h->host_trk[0];
  if (!ht.valid) return fail(h, PFMPE_ERR_STATE, "track_init must precede track_frame_host");
  pfmpe_frame_in in;
  std::memset(&in, 0, sizeof in);
  pf_aff P, pred;
  pf_predict_pose(ht.prev, ht.cur, ht.t_prev, ht.t_cur, time_to_predict, P, pred);   // :234 -> :995-1010
  in.image_points = image_points;
  in.n_image_points = n_image_points;
  in.it_since_initialized = ht.isi;
  pf_aff_to_cm16(ht.cur, in.current_pose);
  pf_aff_to_cm16(pred, in.predicted_pose);
  pf_aff_to_cm16(P, in.prediction_matrix);
  pf_aff_to_cm16(pf_aff_identity(), in.cam_move_inv);
  in.dt = time_to_predict - ht.t_cur;
  in.seed = seed;
  pfmpe_frame_out fo;
  const int rc = pfmpe_shard_frame(hs, n_local, &in, &fo);
  if (rc) return rc;
  std::memset(out, 0, sizeof *out);
  out->status = fo.status; out->flag_fail = (float)fo.status; out->iterations = fo.iterations; out->gn_iterations = -1;
  out->most_likely_pose_idx = fo.most_likely_pose_idx; out->max_count = fo.max_count;
  out->highest_prob = fo.highest_prob; out->weight_sum = fo.weight_sum;
  std::memcpy(out->winner_pose, fo.winner_pose, sizeof out->winner_pose);
  out->n_correspondences = fo.n_correspondences;
  std::memcpy(out->correspondences, fo.correspondences, sizeof out->correspondences);
  if (fo.resampled) {
    pf_aff pose = pf_aff_from_cm16(fo.winner_pose);         // predicted_pose_ = PoseParticle[q] (:687)
    const pf_aff before = pose;
    double c[36];
    out->gn_iterations = pf_optimise_pose(h->K, &h->markers[0], image_points, fo.correspondences, fo.n_correspondences, pose, c);   // :690
    if (ht.isi < 2) ++ht.isi;                               // :2027-2030
    ht.prev = ht.cur; ht.cur = pose;                        // updatePose :2013-2014
    if (time_to_predict - ht.t_cur > 0.001 || time_to_predict < ht.t_cur) { ht.t_prev = ht.t_cur; ht.t_cur = time_to_predict; }
    for (int r = 0; r < 3; ++r)
      for (int q = 0; q < 3; ++q)
        if (fabs(before.a[4 * r + q] - pose.a[4 * r + q]) >= 0.3) out->flag_fail = 1.5f;   // :693-701
    for (int r = 0; r < 36; ++r) ht.cov[r] = c[r];
  } else {
    ht.isi = 0;                                             // :710: the caller must re-initialise
  }
  for (int i = 1; i < n_local; ++i) hs[i]->host_trk[0] = ht;   // co-located shards track the same object
  pf_aff_to_cm16(ht.cur, out->pose);
  for (int r = 0; r < 6; ++r) for (int q = 0; q < 6; ++q) out->cov[r + 6 * q] = ht.cov[6 * r + q];
  return PFMPE_OK;
}

int pfmpe_shard_passes(const pfmpe_handle* h, uint64_t* n)
{
  if (!h || !n) return PFMPE_ERR_INVALID;
  *n = h->shard_passes;
  return PFMPE_OK;
}

// ---- position-keyed checksums of the filter state (full-size parity checks without a host copy) ----------
int pfmpe_state_checksum(pfmpe_handle* h, int tracker, uint64_t out[3])
{
  if (!h || !out) return PFMPE_ERR_INVALID;
  if (tracker < 0 || tracker >= h->cfg.n_trackers) return fail(h, PFMPE_ERR_INVALID, "tracker out of range");
  int rc = fetch_state(h, tracker);
  if (rc) return rc;
  const pf_tracker_state& s = h->h_state[tracker];
  const size_t N = (size_t)h->cfg.n_particles;
  if (!h->cksum) PF_CUDA(h, cudaMalloc((void**)&h->cksum, 3 * sizeof(unsigned long long)));
  PF_CUDA(h, cudaMemsetAsync(h->cksum, 0, 3 * sizeof(unsigned long long), h->stream));
  const float4* T = (s.t_buf == 0 ? h->pose0 : h->pose1) + (size_t)tracker * 3 * h->pstride;
  const float4* src = (s.src_buf == 0 ? h->pose0 : h->pose1) + (size_t)tracker * 3 * h->pstride;
  const int* gidx = (h->sharded && h->lead) ? h->lead : h->idx + (size_t)tracker * N;
  int blocks = (int)((N + PF_BLOCK - 1) / PF_BLOCK);
  if (blocks > h->num_sms * 16) blocks = h->num_sms * 16;
  pf_checksum_kernel<<<blocks, PF_BLOCK, 0, h->stream>>>(T, h->w + (size_t)tracker * N, src, gidx, s.src_mode, (int)N, (int)h->pstride,
                                                          (unsigned long long)(h->sharded ? h->goff : 0), h->cksum);
  PF_CUDA(h, cudaGetLastError());
  h->launches += 1;
  unsigned long long c[3];
  PF_CUDA(h, cudaMemcpyAsync(c, h->cksum, sizeof c, cudaMemcpyDeviceToHost, h->stream));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  for (int i = 0; i < 3; ++i) out[i] = c[i];
  return PFMPE_OK;
}

// ---- streams / timing ---------------------------------------------------------------------
int pfmpe_set_stream(pfmpe_handle* h, void* cuda_stream)
{
  if (!h) return PFMPE_ERR_INVALID;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  if (h->stream) PF_CUDA(h, cudaStreamSynchronize(h->stream));
  if (h->own_stream && h->stream) cudaStreamDestroy(h->stream);
  h->stream = (cudaStream_t)cuda_stream;
  h->own_stream = false;
  return PFMPE_OK;
}

int pfmpe_sync(pfmpe_handle* h)
{
  if (!h) return PFMPE_ERR_INVALID;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  return PFMPE_OK;
}

int pfmpe_last_launch_ms(pfmpe_handle* h, float* ms)
{
  if (!h || !ms) return PFMPE_ERR_INVALID;
  if (!h->timing_valid) return fail(h, PFMPE_ERR_STATE, "nothing has been launched yet");
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaEventSynchronize(h->ev1));
  PF_CUDA(h, cudaEventElapsedTime(ms, h->ev0, h->ev1));
  return PFMPE_OK;
}

int pfmpe_launch_count(const pfmpe_handle* h, uint64_t* n)
{
  if (!h || !n) return PFMPE_ERR_INVALID;
  *n = h->launches;
  return PFMPE_OK;
}

int pfmpe_debug_timeline(pfmpe_handle* h, int enable, uint64_t stamps_ns[16])
{
  if (!h) return PFMPE_ERR_INVALID;
  PF_CUDA(h, cudaSetDevice(h->cfg.device));
  PF_CUDA(h, cudaStreamSynchronize(h->stream));
  if (stamps_ns) PF_CUDA(h, cudaMemcpy(stamps_ns, h->timeline, 16 * sizeof(uint64_t), cudaMemcpyDeviceToHost));
  h->timeline_on = enable != 0;
  return PFMPE_OK;
}

int pfmpe_launch_info(const pfmpe_handle* h, int32_t info[4])
{
  if (!h || !info) return PFMPE_ERR_INVALID;
  for (int i = 0; i < 4; ++i) info[i] = h->info[i];
  return PFMPE_OK;
}

}  // extern "C"
