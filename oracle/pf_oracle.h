/*
 * pf_oracle.h -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * Dependency-free C restatement, in double precision, of the particle-filter
 * tracking core of pf_mpe_lib (reference: /root/reference/pf_mpe_lib/src/
 * pose_estimator.cpp).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference leg may load this library; the product
 * (pf_monocular_pose_estimator_b200/) never links, imports or calls it.
 *
 * PARITY PINNING STATUS
 *   - RNG stream (minstd_rand0 + generate_canonical<double,53> +
 *     uniform_real_distribution<double>): PINNED against this container's
 *     libstdc++ by oracle/check_rng.cpp (bit-exact).
 *   - Everything else (Eigen products / minCoeff visitor order / LDLT):
 *     "parity unpinned" -- the reference has no tests, golden vectors or
 *     fixtures for this path and cannot be compiled here (needs Eigen,
 *     OpenCV C++ headers and ROS, all absent).  The restatement follows the
 *     cited lines; an independent NumPy twin (oracle/pf_oracle_np.py)
 *     cross-checks it.
 *
 * Conventions: 4x4 poses are column-major double[16] (Eigen::Matrix4d
 * layout); K is row-major double[9] (cv::Mat at<double>(i,j) /
 * sensor_msgs/CameraInfo.K order); object points are homogeneous double[4];
 * image points are double[2] undistorted pixels.
 */
#ifndef PF_ORACLE_H_
#define PF_ORACLE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_LED 8

/* ---- RNG: std::default_random_engine == std::minstd_rand0 ------------- */
typedef struct { uint32_t x; } orc_rng;
void     orc_rng_seed(orc_rng *g, uint32_t seed);        /* libstdc++ seeding rule   */
uint32_t orc_rng_next(orc_rng *g);                       /* x <- 16807 x mod 2^31-1  */
double   orc_rng_canonical(orc_rng *g);                  /* generate_canonical<double,53>: 2 draws */
double   orc_rng_uniform(orc_rng *g, double a, double b);/* uniform_real_distribution<double>(a,b) */
/* state after `steps` engine steps from state x (skip-ahead, for tests)  */
uint32_t orc_rng_skip(uint32_t x, uint64_t steps);

/* ---- small 4x4 helpers (column-major) -------------------------------- */
void orc_mat4_mul(const double *A, const double *B, double *C);    /* C = A*B */
void orc_mat4_inv(const double *A, double *Ainv);                   /* general inverse */
void orc_mat4_identity(double *A);

/* ---- project2d  (pose_estimator.cpp:1017-1034) ------------------------ */
void orc_project2d(const double K[9], const double T[16], const double X[4], double uv[2]);

/* ---- LEDDetector::distortPoints, one point (led_detector.cpp:371-414) -- */
void orc_distort_point(const double K[9], const double D5[5], const double in[2], double out[2]);

/* ---- PF parameters (pose_estimator.h:139-155, cfg :25-38) ------------- */
typedef struct {
  double min_angular_noise, max_angular_noise;         /* minAngularNoise / maxAngularNoise       */
  double min_transition_noise, max_transition_noise;   /* minTransitionNoise / maxTransitionNoise */
  double back_projection_pixel_tolerance;              /* gain shape  (:2416)                     */
  double back_projection_pixel_tolerance_PF;           /* accept gate (:2414)                     */
  int    marker_downgrade[5];                          /* bMarkerDowngrade, exactly five (Q5)     */
} orc_params;

/* ---- calculateEstimationProbability (pose_estimator.cpp:2385-2445) ---- */
/* img: D x 2, proj: L x 2.  pairs: out, up to L rows of (LED+1, blob+1).   */
double orc_score(const orc_params *prm, int L, int D, const double *img,
                 const double *proj, unsigned *pairs, int *npairs);

/* ---- one PF update (pose_estimator.cpp:475-733) ----------------------- */
typedef struct {
  int N, L, D;
  const double *obj;     /* L x 4 homogeneous marker positions (object_points_)          */
  const double *img;     /* D x 2 detections (image_points_)                             */
  const double *K;       /* 3x3 row-major                                                */
  const double *E;       /* N x 16: newPoseEstimation (last frame's resampled set)       */
  const double *T_cur;   /* current_pose_                                                */
  const double *T_pred;  /* predicted_pose_ (already camMoveInv * ...)                   */
  const double *P;       /* predictionMatrix                                             */
  const double *C;       /* camMoveInv                                                   */
  unsigned isi;          /* it_since_initialized_ at entry                               */
  double   dt;           /* predicted_time_ - current_time_                              */
  uint32_t seed;         /* injected value of rd() (:476-477) -- the one permitted edit  */
  int      resampler;    /* 0 = literal O(N^2) scan (:668-682); 1 = O(N) first-crossing  */
  int      max_iter;     /* 80 in the reference (:616); exposed for tests                */
  const double *dist;    /* NULL (reference behaviour) or 5 plumb_bob coefficients: option f3, projections go
                            through orc_distort_point and img holds raw blob centres      */
} orc_pf_in;

typedef struct {
  /* caller-allocated arrays */
  double   *T;        /* N x 16 PoseParticle of the kept (best) iteration                */
  double   *w_raw;    /* N raw scores of the kept iteration                              */
  double   *w;        /* N probPart after normalisation (:627-629)                       */
  unsigned *pairs;    /* N x 2*ORC_MAX_LED correspondencesVec of the kept iteration      */
  int      *npairs;   /* N                                                               */
  double   *E_new;    /* N x 16 newPoseEstimation after resampling (unchanged if none)   */
  int      *idx;      /* N Particle_index per resample (-1 if undefined)                 */
  unsigned *count;    /* N counterMeas                                                   */
  /* scalars */
  int      iterations;               /* total do/while iterations executed               */
  int      best_iteration;           /* iteration whose snapshot was kept (-1: none)     */
  double   highest_prob;             /* highestProb                                      */
  double   weight_sum;               /* probPartSum before normalisation                 */
  int      most_likely_particle_idx; /* argmax weight of kept iteration (-1 undefined)   */
  int      most_likely_pose_idx;     /* argmax resample count (-1 if not resampled)      */
  int      accepted;                 /* test at :633                                     */
  int      resampled;                /* :666                                             */
  float    flag_fail;                /* 1 / 4 (1.5 is set later by the caller's jump check) */
  double   winner_pose[16];          /* predicted_pose_ leaving the PF block (before GN) */
  unsigned winner_pairs[2*ORC_MAX_LED];
  int      winner_npairs;
  unsigned isi_out;                  /* it_since_initialized_ leaving the PF block (before optimiseAndUpdatePose) */
  uint64_t rng_steps;                /* engine steps consumed (stream-position check)    */
} orc_pf_out;

int orc_pf_update(const orc_params *prm, const orc_pf_in *in, orc_pf_out *out);

/* stage oracle: stratified resampling only (:668-682) on given normalised weights
 * and given uniforms r[k] in [0,1).  literal!=0 -> O(N^2).  idx[k] = -1 where the
 * reference leaves Particle_index untouched and no previous value exists.        */
void orc_resample(int N, const double *w, const double *r, int literal, int *idx, unsigned *count);

/* ---- host tail: Lie maps, predictPose, Gauss-Newton ------------------- */
void orc_exponential_map(const double twist[6], double T[16]);        /* :2194-2226 */
void orc_logarithm_map(const double T[16], double twist[6]);          /* :2228-2296 */
/* predictPose :995-1010: returns predictionMatrix in P and predicted pose */
void orc_predict_pose(const double prev[16], const double cur[16], double t_prev, double t_cur,
                      double t_pred, double P[16], double pred[16]);
/* optimisePose :1805-2009.  pose: in/out predicted_pose_; cov: 6x6 column-major;
 * returns number of iterations reported in PubData.numIter (or -1 if never converged) */
int  orc_optimise_pose(const double K[9], int L, const double *obj, int D, const double *img,
                       const unsigned *pairs, int npairs, double pose[16], double cov[36]);

/* ---- a whole tracked frame (estimateBodyPose PF branch :223-733 minus detection/ROI) */
typedef struct {
  int N, L;
  double obj[4*ORC_MAX_LED];
  double K[9];
  orc_params prm;
  double current_pose[16], previous_pose[16], predicted_pose[16];
  double current_time, previous_time, predicted_time;
  double pose_covariance[36];
  unsigned isi;
  double *E;      /* N x 16 newPoseEstimation */
  double *T;      /* N x 16 PoseParticle      */
  double *w;      /* N probPart               */
  int resampler;
  /* scratch */
  orc_pf_out pf;
} orc_tracker;

int  orc_tracker_init(orc_tracker *t, int N, int L, const double *obj, const double K[9],
                      const orc_params *prm, const double pose0[16], double t0,
                      const double pose1[16], double t1, int resampler);
void orc_tracker_free(orc_tracker *t);
/* returns Flag_Fail (1, 1.5, 4) */
float orc_tracker_frame(orc_tracker *t, double time_to_predict, int D, const double *img, uint32_t seed);

/* ---- north_star extra, no reference counterpart: parity unpinned */
/* ROI prediction: pose_estimator.cpp:1036-1052 + the min/max loop of led_detector.cpp:221-244 */
void orc_predict_roi(int N, const double *E, const double Cm[16], const double P[16], const double pred[16],
                     const double *markers_xyz, int L, const double K[9], double out[4]);

void orc_weighted_mean_cov(int N, const double *T, const double *w, double mean_pose[16], double cov[36]);

#ifdef __cplusplus
}
#endif
#endif
