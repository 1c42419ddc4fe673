// pf_p3p.cuh -- K9: the correspondence vote of the brute-force (re-)initialisation.
// PoseEstimator::initialise (pose_estimator.cpp:1524-1717) runs Kneip's P3P (p3p.cpp:65-292) for every triple
// of detections x every ordered triple of LEDs -- C(D,3) * L(L-1)(L-2) solves, 8.5e7 at D = 205 -- projects
// the LEDs that were not used through each of the up to four poses and votes (detection, LED) pairs into
// hist_corr whenever a detection lies within back_projection_pixel_tolerance_ of a projected LED.
// Everything here is FP64 like the reference (the quartic is ill-conditioned; FP32 loses real solutions).
//
// Work decomposition: one CTA per pair (i0 < i1) of detections; its items are (i2, LED permutation), one per
// thread and step.  The detections sit in shared memory sorted by x, so a projected LED only looks at the
// detections whose column lies within the tolerance (the vote is exact: every candidate is tested with the
// reference's own rule -- nearest projected LED first, then sqrt(d2) < tol).  Votes go to a shared-memory
// histogram (integer atomics: order-independent, deterministic) that is flushed once per CTA.
#pragma once

#include <cstddef>
#include <cstdint>

#include <cuda_runtime.h>

#define PF_P3P_BLOCK 128
#define PF_P3P_MAX_D 1024
#define PF_P3P_MAX_L 8

struct pf_p3p_d3 { double x, y, z; };
// world part of one P3P problem (intermediate world frame of p3p.cpp:127-148), 16 doubles
struct pf_p3p_world {
  pf_p3p_d3 n1, n2, n3;   // rows of N
  pf_p3p_d3 P1;
  double p1, p2, d12;
  double colinear;        // table entries only: != 0 when the three LEDs are colinear (:78-84), no solution
};

struct pf_p3p_params {
  int D, L;
  double K[9];                 // row-major
  double obj[PF_P3P_MAX_L * 3];
  double tol;                  // back_projection_pixel_tolerance_
  const double* img;           // [D][2] detections, caller order
  double* simg;                // [D][2] detections sorted by x (written by the prologue)
  int* sidx;                   // [D] caller index of sorted detection
  double* vec;                 // [D][3] unit bearing vectors (calculateImageVectors :1072-1085), caller order
  pf_p3p_world* wtab;          // [L(L-1)(L-2)][2] world parts: LED permutation as given / first two swapped
  int nchunk;                  // slices of the permutation list per detection triple (divides the work of small D)
  unsigned* hist;              // [D][L]
  unsigned long long* stats;   // [4] triples passing the pre-filters, P3P solved, solutions projected, solutions voting
};

#define PF_P3P_CHECK_STRIDE 56   // C(8,3): the largest number of result records a correspondence set can need
// one (correspondence set, combination) of checkCorrespondences; same layout as pfmpe_p3p_check_result
struct pf_p3p_check_rec {
  int rc, valid, best, reserved;
  double squared_error;
  double pose[16];
};

// host-side launchers (pf_p3p.cu, compiled with -fmad=false: the reference is an x86-64 build without FMA
// contraction, and the duplicate-solution test of :1659-1663 compares doubles for exact equality)
size_t pf_p3p_smem_bytes(int D, int L);
size_t pf_p3p_roots_bytes_per_thread(int L, int nchunk);
// prologue + vote; with a roots buffer (rbuf, rbuf_bytes) the vote runs as roots kernel + back-projection kernel per slab
cudaError_t pf_p3p_launch_hist(const pf_p3p_params& p, cudaStream_t stream, double2* rbuf = nullptr, size_t rbuf_bytes = 0);
cudaError_t pf_p3p_launch_check(const pf_p3p_params& p, int M, int stride, const unsigned* corr, const int* counts,
                                double certainty_threshold, pf_p3p_check_rec* out, cudaStream_t stream);
cudaError_t pf_p3p_launch_solve(int n, const double* fv, const double* wp, double* sol, int* rc, cudaStream_t stream);
