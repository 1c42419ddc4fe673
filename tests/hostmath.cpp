// hostmath.cpp -- TEST ARTEFACT ONLY (built by tests/, never shipped, never linked by the product).
// Compiles the device math headers (pf_math.cuh / pf_tail.cuh, PF_HD functions) with g++ and
// replays one frame of the kernel's algorithm sequentially, so that the FP32 arithmetic, the
// fixed-point weight sum, the (sum, running-max) scan, K(m) thresholds and the source-index
// search can be checked against the oracle on a machine without a GPU.
#include <cstdint>
#include <cstring>
#include <vector>

#include "../pf_monocular_pose_estimator_b200/csrc/pf_tail.cuh"

namespace {

template <int L>
float score_particle(const pf_pose& T, const float* K9f, const float* obj, const pf_f2* blobs, int D,
                     const pf_score_cfg& sc, unsigned* pairs, int* npairs)
{
  float u[L], v[L];
  const bool kstd = pf_k_is_std(K9f);
  for (int j = 0; j < L; ++j) pf_project(K9f, kstd, T, obj[3 * j], obj[3 * j + 1], obj[3 * j + 2], u[j], v[j]);
  // the device computes every particle's weight with the unsorted form (D >= L) and only the winner's correspondences
  // with the sorted greedy loop: same here
  const float wgt = pf_score<L, false>(blobs, D, u, v, sc, nullptr, nullptr);
  (void)pf_score<L, true>(blobs, D, u, v, sc, pairs, npairs);
  return wgt;
}

template <int L>
int frame(int N, int D, const double* markers, const double* K9, const double* blobs_d, const double* E,
          const double* cur, const double* pred, const double* P, const double* C, unsigned isi, double dt,
          uint32_t seed, const double* prm, const int* dg, float* T_out, float* w_out, int* kend_out, int* idx_out,
          int* sc_out, long long* sum_fx, unsigned* pairs_out)
{
  std::vector<uint32_t> tab(PF_RNG_TAB_WORDS);
  pf_build_rng_table(tab.data());
  pf_frame_const fc;
  pf_make_frame_const(fc, pf_aff_from_cm16(cur), pf_aff_from_cm16(pred), pf_aff_from_cm16(P), pf_aff_from_cm16(C), isi,
                      dt, seed, D, prm[0], prm[1], prm[2], prm[3]);
  pf_score_cfg sc;
  sc.tol = (float)prm[4]; sc.inv_tol = (float)(1.0 / prm[4]); sc.tol_pf = (float)prm[5];
  sc.dg_mask = 0;
  for (int i = 0; i < 5; ++i) if (dg[i]) sc.dg_mask |= 1u << i;
  pf_score_cfg_finish(sc, L);
  float K9f[9], obj[PF_MAX_LED * 3];
  for (int i = 0; i < 9; ++i) K9f[i] = (float)K9[i];
  for (int i = 0; i < 3 * L; ++i) obj[i] = (float)markers[i];
  std::vector<pf_f2> blobs(D > 0 ? D : 1);
  for (int i = 0; i < D; ++i) { blobs[i].x = (float)blobs_d[2 * i]; blobs[i].y = (float)blobs_d[2 * i + 1]; }
  std::vector<pf_pose> Ef(N), T(N);
  for (int n = 0; n < N; ++n) Ef[n] = pf_aff_to_f32(pf_aff_from_cm16(E + 16 * (size_t)n));
  pf_pose Pm, Cm;
  for (int i = 0; i < 12; ++i) { Pm.p[i] = fc.P[i]; Cm.p[i] = fc.C[i]; }
  std::vector<float> w(N);

  auto pass = [&](int it, float& maxw, int& argmax, int& argmin, long long& S) {
    const bool use_cam = fc.isi > 1, use_pred = use_cam && (it % 10) != 0;
    const bool cam_mul = use_cam && !(fc.flags & PF_FC_CAM_IDENTITY);
    const int ang_mode = (fc.flags & PF_FC_TINY_ANG) ? 2 : ((fc.flags & PF_FC_SMALL_ANG) ? 1 : 0);
    const float scale = 1.0f + 0.025f * (float)(it / 10);
    const uint64_t step0 = 12ull * (uint64_t)(N - 2) * (uint64_t)it;
    S = 0;
    float minw = 0;
    for (int n = 0; n < N; ++n) {
      pf_pose Tn;
      if (n == 0) for (int i = 0; i < 12; ++i) Tn.p[i] = fc.Tcur[i];
      else if (n == 1) for (int i = 0; i < 12; ++i) Tn.p[i] = fc.Tpred[i];
      else {
        pf_pose B = Ef[n];
        if (cam_mul) B = pf_affine_mul(Cm, B);
        if (use_pred) B = pf_affine_mul(B, Pm);
        uint32_t x = pf_rng_skip(fc.seed_state, step0 + 12ull * (uint64_t)(n - 2), tab.data());
        Tn = pf_propagate(B, x, fc.noise, scale, ang_mode);
      }
      T[n] = Tn;
      unsigned pr[2 * PF_MAX_LED]; int np;
      w[n] = score_particle<L>(Tn, K9f, obj, blobs.data(), D, sc, pr, &np);
      if (n == 0 || w[n] > maxw) { maxw = w[n]; argmax = n; }
      if (n == 0 || w[n] < minw) { minw = w[n]; argmin = n; }
      S += pf_w_to_fx(w[n]);
    }
  };

  const float thr_exit = (float)(L * (D < 5 ? D : 5));
  float highest = 0;
  int best_it = -1, it = 0, k_arg = 0, k_argmin = 0, l_arg = 0, l_argmin = 0;
  long long k_S = 0, l_S = 0;
  for (;;) {
    float m; pass(it, m, l_arg, l_argmin, l_S);
    ++it;
    if (m > highest) { highest = m; best_it = it - 1; k_arg = l_arg; k_argmin = l_argmin; k_S = l_S; }
    if (!(it < 80 && m < thr_exit)) break;
  }
  const int iters = it;
  if (best_it < 0) { k_arg = l_arg; k_argmin = l_argmin; k_S = l_S; }
  else if (best_it != iters - 1) { float m; int a, b; long long s; pass(best_it, m, a, b, s); }
  const bool accept = (k_S != 0) && (highest > (float)(L * (D < 3 ? D : 3)));
  const bool resampled = accept && fc.isi > 0;
  int q = k_S < 0 ? k_argmin : k_arg;
  unsigned maxcount = 0;
  if (resampled) {
    const long long sg = k_S < 0 ? -1 : 1;
    const uint32_t x_base = pf_rng_skip(fc.seed_state, 12ull * (uint64_t)(N - 2) * (uint64_t)iters, tab.data());
    const pf_resample_cfg rc = pf_make_resample_cfg(x_base, N, k_S);
    pf_sm pre = pf_sm_identity();
    int kprev = 0;
    for (int i = 0; i < N; ++i) {
      pf_sm e; e.s = sg * pf_w_to_fx(w[i]); e.m = e.s;
      pre = pf_sm_combine(pre, e);
      kend_out[i] = pf_count_leq(pre.m, rc, tab.data());
      unsigned cnt = (unsigned)(kend_out[i] - kprev);
      if (cnt > maxcount) { maxcount = cnt; q = i; }
      kprev = kend_out[i];
    }
    const int ktot = kend_out[N - 1];
    for (int n = 0; n < N; ++n) {
      int k = n < ktot ? n : ktot - 1;
      if (k < 0) { idx_out[n] = n; continue; }
      int a = 0;
      while (kend_out[a] <= k) ++a;
      idx_out[n] = a;
    }
  }
  for (int n = 0; n < N; ++n) { std::memcpy(T_out + 12 * (size_t)n, T[n].p, sizeof(float) * 12); w_out[n] = w[n]; }
  int np = 0;
  (void)score_particle<L>(T[q], K9f, obj, blobs.data(), D, sc, pairs_out, &np);
  sc_out[0] = iters; sc_out[1] = best_it; sc_out[2] = accept; sc_out[3] = resampled; sc_out[4] = k_arg;
  sc_out[5] = q; sc_out[6] = (int)maxcount; sc_out[7] = np;
  *sum_fx = k_S;
  ((float*)sc_out)[8] = highest;
  return 0;
}

}  // namespace

extern "C" {

int hm_frame(int N, int L, int D, const double* markers, const double* K9, const double* blobs, const double* E,
             const double* cur, const double* pred, const double* P, const double* C, unsigned isi, double dt,
             uint32_t seed, const double* prm, const int* dg, float* T_out, float* w_out, int* kend_out, int* idx_out,
             int* sc_out, long long* sum_fx, unsigned* pairs_out)
{
#define HM_CASE(LL) case LL: return frame<LL>(N, D, markers, K9, blobs, E, cur, pred, P, C, isi, dt, seed, prm, dg, T_out, w_out, kend_out, idx_out, sc_out, sum_fx, pairs_out);
  switch (L) { HM_CASE(3) HM_CASE(4) HM_CASE(5) default: return -1; }
}

// resampling stage alone on given raw FP32 weights: kend + source indices
int hm_resample(int N, const float* w, uint32_t seed, int iters, int* kend_out, int* idx_out, long long* sum_fx)
{
  std::vector<uint32_t> tab(PF_RNG_TAB_WORDS);
  pf_build_rng_table(tab.data());
  long long S = 0;
  for (int i = 0; i < N; ++i) S += pf_w_to_fx(w[i]);
  *sum_fx = S;
  if (S == 0) return 1;
  uint32_t x0 = seed % PF_RNG_M; if (!x0) x0 = 1;
  const long long sg = S < 0 ? -1 : 1;
  const uint32_t x_base = pf_rng_skip(x0, 12ull * (uint64_t)(N - 2) * (uint64_t)iters, tab.data());
  const pf_resample_cfg rc = pf_make_resample_cfg(x_base, N, S);
  pf_sm pre = pf_sm_identity();
  for (int i = 0; i < N; ++i) {
    pf_sm e; e.s = sg * pf_w_to_fx(w[i]); e.m = e.s;
    pre = pf_sm_combine(pre, e);
    kend_out[i] = pf_count_leq(pre.m, rc, tab.data());
  }
  const int ktot = kend_out[N - 1];
  int a = 0;
  for (int n = 0; n < N; ++n) {
    int k = n < ktot ? n : ktot - 1;
    if (k < 0) { idx_out[n] = n; continue; }
    while (kend_out[a] <= k) ++a;
    idx_out[n] = a;
  }
  return 0;
}

uint32_t hm_rng_skip(uint32_t x, uint64_t steps)
{
  std::vector<uint32_t> tab(PF_RNG_TAB_WORDS);
  pf_build_rng_table(tab.data());
  return pf_rng_skip(x, steps, tab.data());
}

double hm_canonical_f64(uint32_t g1, uint32_t g2) { return pf_canonical_f64(g1, g2); }
float hm_canonical_f32(uint32_t g1, uint32_t g2) { return pf_canonical_f32(g1, g2); }

}  // extern "C"
