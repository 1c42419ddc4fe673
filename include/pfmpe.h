/*
 * pfmpe.h -- C ABI of the B200-native particle-filter tracking core (libpfmpe.so).
 *
 * Drop-in boundary for the PF block that pf_mpe_lib inlines in
 * PoseEstimator::estimateBodyPose (/root/reference/pf_mpe_lib/src/pose_estimator.cpp:471-735):
 * everything between setImagePoints() (:469) and optimiseAndUpdatePose() (:690) is replaced by
 * pfmpe_update(); the reference has no FFI for this path, so each entry point cites the
 * PoseEstimator member / source lines it stands in for (header: pf_mpe_lib/include/pf_mpe_lib/
 * pose_estimator.h).  INTEGRATION.md shows the shim a maintainer adds on the reference side.
 *
 * Conventions: C linkage, POD structs, caller owns every host buffer, the handle owns device
 * memory and its stream.  Every function returns 0 on success or a negative pfmpe_status; it
 * never throws and never falls back to a CPU path (a missing/failed CUDA device is an error).
 * One handle is single-threaded like the reference (callers serialise); handles are independent.
 * Matrices: 4x4 poses are column-major double[16] (Eigen::Matrix4d memory layout); K is row-major
 * double[9] (cv::Mat / sensor_msgs::CameraInfo.K order); 6x6 covariances are column-major.
 */
#ifndef PFMPE_H_
#define PFMPE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PFMPE_MAX_LED 8
#define PFMPE_VERSION 1

typedef enum {
  PFMPE_OK = 0,
  PFMPE_ERR_INVALID = -1,     /* bad argument                                       */
  PFMPE_ERR_CUDA = -2,        /* a CUDA call failed; see pfmpe_last_error()          */
  PFMPE_ERR_NO_DEVICE = -3,   /* no usable CUDA device (there is no CPU fallback)    */
  PFMPE_ERR_STATE = -4,       /* call sequence error (e.g. update before set_particles) */
  PFMPE_ERR_UNSUPPORTED = -5, /* configuration outside the supported range           */
  PFMPE_ERR_COMM = -6         /* multi-GPU communicator error                        */
} pfmpe_status;

typedef struct pfmpe_handle pfmpe_handle;

/* flags */
#define PFMPE_FLAG_MEAN_COV 1u /* also reduce to a weighted mean pose + covariance (north_star extra) */
#define PFMPE_FLAG_DISTORT 2u  /* project through K AND the plumb_bob coefficients of pfmpe_set_camera (model of
                                * LEDDetector::distortPoints, led_detector.cpp:371-414): image_points are then the raw,
                                * still-distorted blob centres.  Not what the reference PF does (it scores undistorted
                                * points with K only); per-frame API only. */

typedef struct {
  int32_t n_particles;  /* N_Particle (pose_estimator.h:139)                                  */
  int32_t n_trackers;   /* tracked objects sharing the handle (object_points_Vec.size())      */
  int32_t n_leds;       /* LEDs per object, 1..PFMPE_MAX_LED (object_points_.size())          */
  int32_t max_blobs;    /* capacity for detections per frame (image_points_.size())           */
  int32_t device;       /* CUDA ordinal                                                       */
  uint32_t flags;
  int32_t reserved[4];
} pfmpe_config;

/* PF parameters: public fields of PoseEstimator (pose_estimator.h:139-155), set by
 * MPENode::dynamicParametersCallback (pf_mpe/src/monocular_pose_estimator.cpp:480-527). */
typedef struct {
  double min_angular_noise, max_angular_noise;           /* minAngularNoise / maxAngularNoise        */
  double min_transition_noise, max_transition_noise;     /* minTransitionNoise / maxTransitionNoise  */
  double back_projection_pixel_tolerance;                /* setBackProjectionPixelTolerance(), :2416 */
  double back_projection_pixel_tolerance_PF;             /* back_projection_pixel_tolerance_PF, :2414 */
  int32_t marker_downgrade[5];                           /* bMarkerDowngrade (bMarkerNr1..5), :2431  */
  int32_t reserved;
} pfmpe_pf_params;

/* Inputs of one PF update = the members the PF block reads at :475-616. */
typedef struct {
  const double* image_points;    /* image_points_ (setImagePoints, :469): n x 2 undistorted pixels */
  int32_t n_image_points;        /* numLED = detected_led_positions.size() (:449)                  */
  uint32_t it_since_initialized; /* it_since_initialized_ (:488,:555)                              */
  double current_pose[16];       /* current_pose_ (:547)                                           */
  double predicted_pose[16];     /* predicted_pose_ = camMoveInv * predicted_pose_ (:395,:551)     */
  double prediction_matrix[16];  /* predictionMatrix = predictPose() (:234)                        */
  double cam_move_inv[16];       /* camMoveInv (:393)                                              */
  double dt;                     /* predicted_time_ - current_time_ (:499)                         */
  uint32_t seed;                 /* value of rd() at :476-477 (the reference draws it per frame)   */
  uint32_t reserved;
} pfmpe_frame_in;

/* Outputs of one PF update = what the PF block leaves behind at :619-719. */
typedef struct {
  int32_t status;                   /* Flag_Fail written by the PF block: 1 (:635) or 4 (:711)     */
  int32_t iterations;               /* do/while iterations executed (:535-616)                    */
  int32_t best_iteration;           /* iteration kept by the snapshot logic (:608-624), -1 none   */
  int32_t resampled;                /* 1 if :666-682 ran                                          */
  double highest_prob;              /* highestProb (:519,:610)                                    */
  double weight_sum;                /* probPartSum before normalisation (:627)                    */
  int32_t most_likely_particle_idx; /* mostLikelyParticleIdx (:610 / :714)                        */
  int32_t most_likely_pose_idx;     /* mostLikelyPoseIdx (:686), -1 if not resampled              */
  uint32_t max_count;               /* counterMeas.maxCoeff() (:686)                              */
  int32_t n_correspondences;        /* correspondences_.rows() (:688)                             */
  uint32_t correspondences[2 * PFMPE_MAX_LED]; /* rows (LED, detection), 1-based (:2417-2418)     */
  double winner_pose[16];           /* predicted_pose_ leaving the block (:687 / :716)            */
  double mean_pose[16];             /* PFMPE_FLAG_MEAN_COV only                                    */
  double mean_cov[36];              /* PFMPE_FLAG_MEAN_COV only                                    */
} pfmpe_frame_out;

/* ---- lifetime ---------------------------------------------------------------------------- */
int pfmpe_create(const pfmpe_config* cfg, pfmpe_handle** out);
int pfmpe_destroy(pfmpe_handle* h);
const char* pfmpe_strerror(int status);
const char* pfmpe_last_error(const pfmpe_handle* h); /* detail of the last failure on this handle */
int pfmpe_version(void);

/* ---- configuration ----------------------------------------------------------------------- */
/* setMarkerPositions (pose_estimator.h:445, pose_estimator.cpp:56-62): xyz = n_leds x 3 metres;
 * tracker = -1 sets every tracker. */
int pfmpe_set_markers(pfmpe_handle* h, int tracker, const double* xyz, int n_leds);
/* camera_matrix_K_ / camera_distortion_coeffs_ (pose_estimator.h:121-122).  D may be NULL; the PF
 * projects with K only (project2d, :1017-1034) -- detections arrive already undistorted. */
int pfmpe_set_camera(pfmpe_handle* h, const double K[9], const double* D5);
int pfmpe_set_params(pfmpe_handle* h, const pfmpe_pf_params* prm);
/* PoseParticle_Vec[t] = newPoseEstimation_Vec[t] = poses (:190-191 after initialise()):
 * n_particles x 16 column-major doubles. */
int pfmpe_set_particles(pfmpe_handle* h, int tracker, const double* poses, int n);

/* ---- the hot path ------------------------------------------------------------------------- */
/* One PF update of one tracker (:475-719).  Synchronous: outputs are valid on return. */
int pfmpe_update(pfmpe_handle* h, int tracker, const pfmpe_frame_in* in, pfmpe_frame_out* out);
/* One PF update of every tracker of the handle (the `for objectNumber` loop at :89, batched). */
int pfmpe_update_batch(pfmpe_handle* h, const pfmpe_frame_in* in, pfmpe_frame_out* out);

/* ---- state getters (getPoseParticles :917, getResampledParticles :923, probPart_Vec :733) -- */
/* which = 0: PoseParticle (propagated, kept iteration); 1: newPoseEstimation (resampled). */
int pfmpe_get_particles(pfmpe_handle* h, int tracker, int which, double* poses /* n x 16 */);
/* normalised != 0: probPart after :627-629; else the raw scores. */
int pfmpe_get_weights(pfmpe_handle* h, int tracker, int normalised, double* w /* n */);
/* Particle_index per resample slot (:678-681) and counterMeas (:524,:678); valid after a resampled frame. */
int pfmpe_get_resample_indices(pfmpe_handle* h, int tracker, int32_t* idx /* n */);
int pfmpe_get_resample_counts(pfmpe_handle* h, int tracker, uint32_t* count /* n */);

/* ---- host tail (FP64), used by the reference-side shim between two updates ----------------- */
/* optimisePose (:1805-2009): pose in/out, cov 6x6; *iters = PubData.numIter or -1.  correspondences: pairs
 * (LED, detection), 1-based; a pair outside 1..n_leds / 1..n_image_points is PFMPE_ERR_INVALID. */
int pfmpe_optimise_pose(const double K[9], const double* markers_xyz, int n_leds, const double* image_points,
                        int n_image_points, const uint32_t* correspondences, int n_correspondences,
                        double pose[16], double cov[36], int* iters);
/* predictPose (:995-1010) */
int pfmpe_predict_pose(const double previous_pose[16], const double current_pose[16], double previous_time,
                       double current_time, double time_to_predict, double prediction_matrix[16],
                       double predicted_pose[16]);
int pfmpe_exponential_map(const double twist[6], double pose[16]);  /* :2194-2226 */
int pfmpe_logarithm_map(const double pose[16], double twist[6]);    /* :2228-2296 */

/* ---- device-resident tracking (whole sequences without a host round trip) ------------------ */
/* Mirrors the PF branch of estimateBodyPose (:223-733) minus detection/ROI: predictPose, PF update,
 * optimiseAndUpdatePose and the jump check all run on the device.
 * State after "initialise": previous_pose_/current_pose_ and their times, it_since_initialized_ = 2,
 * every particle = pose1 (:106-109,:190-191). */
int pfmpe_track_init(pfmpe_handle* h, int tracker, const double pose0[16], double t0, const double pose1[16],
                     double t1);
/* The same for a filter whose particles were seeded individually (pfmpe_set_particles first): what initialise() +
 * optimiseAndUpdatePose leave behind at :166-186 -- previous_pose_ / current_pose_ and their times,
 * it_since_initialized_ (1 right after an initialisation), PoseParticle = newPoseEstimation = the P3P seeds.
 * previous_time == current_time is rejected (PFMPE_ERR_INVALID; also t0 == t1 in pfmpe_track_init): the next
 * predictPose divides by their difference (:1002). */
int pfmpe_track_set_state(pfmpe_handle* h, int tracker, const double previous_pose[16], double previous_time,
                          const double current_pose[16], double current_time, uint32_t it_since_initialized);
/* Upload a sequence: image_points [n_trackers][n_frames][max_blobs][2], counts [n_trackers][n_frames],
 * times [n_frames] (time_to_predict per frame), seeds [n_trackers][n_frames]. */
int pfmpe_track_upload(pfmpe_handle* h, int n_frames, const double* image_points, const int32_t* counts,
                       const double* times, const uint32_t* seeds);
/* Same, from device pointers already resident in HBM (no copy is made; buffers must outlive the runs). */
int pfmpe_track_attach(pfmpe_handle* h, int n_frames, const void* d_image_points, const void* d_counts,
                       const void* d_times, const void* d_seeds);
/* Run frames [first, first + n) of the uploaded sequence for all trackers. Asynchronous on the handle's
 * stream; pfmpe_sync() or a getter completes it. */
int pfmpe_track_run(pfmpe_handle* h, int first_frame, int n_frames);

/* Re-run the PF updates (:475-689) of frames that pfmpe_track_run has already processed, reading each
 * frame's constants (current/predicted pose, predictionMatrix, noise ranges, seed) from the log the tracked
 * run left in HBM: the hot path alone, chained frame to frame on the device, without the FP64 tail.  After
 * pfmpe_track_init has restored the same starting particles the results are identical to the tracked run's. */
int pfmpe_track_replay(pfmpe_handle* h, int first_frame, int n_frames);

typedef struct {
  int32_t status;       /* 0 not run (track lost earlier), 1 accepted, 4 rejected                   */
  float flag_fail;      /* PubData.Flag_Fail: 1, 1.5 (:699) or 4                                    */
  int32_t iterations;
  int32_t gn_iterations; /* PubData.numIter (:1883)                                                  */
  int32_t most_likely_pose_idx;
  uint32_t max_count;
  double highest_prob, weight_sum;
  double pose[16];      /* current_pose_ after updatePose (getPredictedPose, :905)                  */
  double cov[36];       /* pose_covariance_ (getPoseCovariance, :911)                               */
  double winner_pose[16];
  int32_t n_correspondences;
  uint32_t correspondences[2 * PFMPE_MAX_LED];
  int32_t reserved;
} pfmpe_track_result;
/* One tracked frame with HOST buffers: the PF branch of PoseEstimator::estimateBodyPose (:223-733) minus
 * detection/ROI -- predictPose (host, FP64), PF update (device: detections and poses go host->device, the result
 * record comes back), optimiseAndUpdatePose + jump check (host, FP64).  The tracked object (previous/current
 * pose, times, it_since_initialized_) is the one pfmpe_track_init set up; image_points: n x 2 undistorted pixels.
 * The batch form does the same for every tracker of the handle with one device launch:
 * image_points [n_trackers][max_blobs][2], counts / seeds [n_trackers]. */
int pfmpe_track_frame_host(pfmpe_handle* h, int tracker, double time_to_predict, const double* image_points,
                           int32_t n_image_points, uint32_t seed, pfmpe_track_result* out);
int pfmpe_track_frame_host_batch(pfmpe_handle* h, double time_to_predict, const double* image_points,
                                 const int32_t* counts, const uint32_t* seeds, pfmpe_track_result* out);

/* The same per-frame call in a loop over a sequence held in HOST memory (layouts of pfmpe_track_upload; out
 * [n_trackers][n_frames]); every frame makes its own host->device / device->host round trip and host tail.
 * frame_seconds (optional, n_frames): wall time of each frame's call. */
int pfmpe_track_sequence_host(pfmpe_handle* h, int n_frames, const double* times, const double* image_points,
                              const int32_t* counts, const uint32_t* seeds, pfmpe_track_result* out,
                              double* frame_seconds);

/* The same tracked frame for EVERY tracker of the handle with HOST buffers in and out, but with the FP64 tail on the
 * device: predictPose (:234 -> :995-1010), the PF update (:475-689), optimiseAndUpdatePose (:690,
 * :1805-2009, :2011-2035) and the jump check (:693-701) run in one launch of the track-mode kernel; per call the
 * detections go host -> device through pinned staging and the result records come back.  For batches of tracked objects
 * (the `for objectNumber` loop at :89 with thousands of objects), where the host-tail form spends its time in the
 * Gauss-Newton of every object on the host cores.  Poses agree with the host-tail form to ~1e-9 (same algorithm, FP64,
 * different operation order).  The device runs at most 64 Gauss-Newton iterations per object (PFMPE_GN_CAP in the
 * environment; 0 = the reference's full 500): the ~1 % of objects still iterating then (the median is 8, a few never
 * converge and run all 500) would hold the whole launch for their serial chain, so their records are packed into pinned
 * memory first, the host cores redo exactly those optimisations (pf_optimise_pose, full :1805-2009 semantics) under the
 * copy of the full record array, and one small kernel patches the device state -- the results are those of the uncapped
 * form.  PFMPE_TRACE=1 prints the per-phase times of this call to stderr every 20 frames.  Layouts as pfmpe_track_frame_host_batch / pfmpe_track_sequence_host; the tracked objects'
 * state (pfmpe_track_init) is shared with the host-tail calls and kept in step. */
int pfmpe_track_frame_device_batch(pfmpe_handle* h, double time_to_predict, const double* image_points,
                                   const int32_t* counts, const uint32_t* seeds, pfmpe_track_result* out);
int pfmpe_track_sequence_device(pfmpe_handle* h, int n_frames, const double* times, const double* image_points,
                                const int32_t* counts, const uint32_t* seeds, pfmpe_track_result* out,
                                double* frame_seconds);

/* results of frames [first, first+n) of one tracker */
int pfmpe_track_results(pfmpe_handle* h, int tracker, int first_frame, int n_frames, pfmpe_track_result* out);

/* ---- ROI prediction for the LED detector ------------------------------------------
 * predictMarkerPositionsInImage (pose_estimator.cpp:1036-1052, called :396) + the extreme-value loop of
 * LEDDetector::determineROI (led_detector.cpp:221-244): every LED through every resampled particle as
 * cam_move_inv * newPoseEstimation[j] * prediction_matrix, and through predicted_pose;
 * roi = {x_min, y_min, x_max, y_max} with the reference's start values (+INF, +INF, 0, 0).  The caller
 * passes the two corner points on to determineROI (distortion, border, clamping: :246-260) instead of
 * the N*L + L predicted pixel positions.  FP32 projection (1e-5 relative of the reference's doubles). */
int pfmpe_predict_roi(pfmpe_handle* h, int tracker, const double cam_move_inv[16], const double prediction_matrix[16],
                      const double predicted_pose[16], double roi[4]);

/* ---- correspondence vote of the brute-force (re-)initialisation --------------------
 * The histogram loop of PoseEstimator::initialise (pose_estimator.cpp:1524-1717): Kneip's P3P
 * (P3P::computePoses / solveQuartic, p3p.cpp:65-292) for every triple of detections x every ordered
 * triple of LEDs, back-projection of the LEDs that were not used through each finite solution
 * (project2d :1017-1034, calculateMinDistancesAndPairs :2093-2137) and the votes of :1692-1712, with the
 * pre-filters of :1556-1581 and calculateImageVectors (:1072-1085) for the bearing vectors.  FP64 on the
 * device, like the reference.  image_points: n x 2 undistorted pixels (3 <= n <= 1024); markers, K and
 * back_projection_pixel_tolerance are the handle's.  hist: n x n_leds, row-major = hist_corr(detection, LED);
 * the caller feeds it to correspondencesFromHistogram (:1134-1288) / checkCorrespondences (:1312-1501)
 * unchanged.  Counts are integers: identical to the reference except where a vote sits within rounding of
 * the tolerance or a conjugate root pair collapses within rounding.
 * Device memory: a few KB of tables per handle, plus -- for votes of 2e6 P3P problems and more (about 60 detections with
 * 5 LEDs) -- a buffer of 32 B per problem for the roots between the vote's two kernels, allocated on first use, kept with
 * the handle and capped at PFMPE_P3P_ROOTS_MB (environment, default 3072; larger votes run in slabs, 0 = never: one kernel). */
typedef struct {
  uint64_t triples;             /* detection triples that pass the pre-filters (:1561-1581)            */
  uint64_t p3p_solved;          /* computePoses calls that returned 0                                  */
  uint64_t solutions_projected; /* finite solutions not dropped by the duplicate test (:1659-1670)     */
  uint64_t solutions_voted;     /* solutions with at least one detection inside the tolerance (:1692)  */
  float device_ms;              /* device time of the vote (CUDA events on the handle's stream)        */
  int32_t reserved;
} pfmpe_p3p_stats;
int pfmpe_p3p_histogram(pfmpe_handle* h, int tracker, const double* image_points, int32_t n_image_points,
                        uint32_t* hist, pfmpe_p3p_stats* stats);
/* The same with image_points already resident in device memory (no host->device copy of the detections;
 * hist still comes back to the host: it is the result). */
int pfmpe_p3p_histogram_dev(pfmpe_handle* h, int tracker, const double* d_image_points, int32_t n_image_points,
                            uint32_t* hist, pfmpe_p3p_stats* stats);

/* The same vote with NO allocation (no reference counterpart: added for callers
 * that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The
 * kernels' scratch (device copy of the detections, copy sorted by x, bearing vectors, world table of the LED
 * permutations, counters, histogram) lives in a file-scope device array of 128 KiB that every GPU loading the library
 * holds until exit; the call makes no cudaMalloc / cudaFree.  Limit: n_image_points <= PFMPE_P3P_STATIC_MAX_POINTS,
 * beyond it PFMPE_ERR_INVALID (use pfmpe_p3p_histogram).  image_points and hist are HOST buffers as for
 * pfmpe_p3p_histogram; results are identical to it.  All calls on one device share the array, which keeps the last
 * call's contents (each call's prologue kernel rewrites its tables and zeroes histogram and counters): calls that
 * overlap on the GPU -- two handles or streams at once -- race; serialise them. */
#define PFMPE_P3P_STATIC_MAX_POINTS 256
int pfmpe_p3p_histogram_static(pfmpe_handle* h, int tracker, const double* image_points, int32_t n_image_points,
                               uint32_t* hist, pfmpe_p3p_stats* stats);

/* The per-combination body of PoseEstimator::checkCorrespondences (pose_estimator.cpp:1340-1445), batched over
 * n_sets candidate correspondence sets (the entries of correspondencesFromHistogram's list, :1733-1751): for every
 * combination of three pairs of a set (Combinations::combinationsNoReplacement order = lexicographic) P3P on the
 * three, back-projection of the set's other LEDs through each finite solution, squared reprojection error and
 * certainty against their detections (calculateSquaredReprojectionErrorAndCertainty, :1087-1132, tolerance =
 * back_projection_pixel_tolerance), and the best solution that reaches certainty_threshold (:1412-1420).
 * correspondences: n_sets rows of 2*PFMPE_MAX_LED uint32 = pairs (LED, detection), 1-based; counts[s] = pairs in
 * row s (3..PFMPE_MAX_LED).  results: n_sets x stride records, record [s][c] for combination c (those beyond
 * C(counts[s], 3) are zeroed); stride >= C(max counts, 3), at most PFMPE_P3P_CHECK_STRIDE = C(8, 3).  The caller counts the valid ones against
 * valid_correspondence_threshold_ (:1469), averages pose * object_points (:1447-1451) and seeds PoseParticle with
 * the poses (:1432-1438). */
#define PFMPE_P3P_CHECK_STRIDE 56
typedef struct {
  int32_t rc;           /* computePoses' return value (0, -1 colinear LEDs), -2 = a pair outside 1..n_leds / 1..n   */
  int32_t valid;        /* valid_correspondence_found (:1412)                                                    */
  int32_t best;         /* smallest_error_idx (:1419), -1 if none                                                */
  int32_t reserved;
  double squared_error; /* min_squared_error (:1418)                                                              */
  double pose[16];      /* temp_solution.inverse() (:1437, :1449), column-major                                  */
} pfmpe_p3p_check_result;
int pfmpe_p3p_check(pfmpe_handle* h, int tracker, const double* image_points, int32_t n_image_points,
                    const uint32_t* correspondences, const int32_t* counts, int32_t n_sets, int32_t stride,
                    double certainty_threshold, pfmpe_p3p_check_result* results);

/* Batched P3P::computePoses (p3p.cpp:65-236), one problem per device thread -- what checkCorrespondences
 * (:1382) and P3P_short (:2506-) call per correspondence triple.  feature_vectors, world_points: n x 9,
 * each a column-major 3x3 (columns = the three unit bearing vectors / world points, Eigen::Matrix3d
 * layout); solutions: n x 4 x 12, each a column-major 3x4 [R | C]; rc: n, 0 or -1 (colinear world points,
 * solutions untouched). */
int pfmpe_p3p_solve(pfmpe_handle* h, int32_t n, const double* feature_vectors, const double* world_points,
                    double* solutions, int32_t* rc);

/* ---- one filter sharded over several GPUs (BASELINE configs[4]) ----------------------
 * One handle per rank owns the contiguous global particles [global_offset, global_offset +
 * n_particles) of ONE filter of n_global particles (n_trackers must be 1).  The reference loop
 * (pose_estimator.cpp:535-616 do/while, :621-633 sum / accept, :666-686 resampling) is split at its
 * global reductions; every stage below is one asynchronous launch on the handle's stream and the
 * exchange steps between them are the caller's (pf_monocular_pose_estimator_b200/sharded.py uses
 * torch.distributed: all_gather of the partials = the "weight-sum allreduce", all_to_all of rows
 * = the "global resample").  RNG, thresholds and tie rules use GLOBAL particle / slot numbers, so
 * the sharded filter reproduces the single-GPU filter bit for bit.
 *
 * d_partial points to 8 int64 in DEVICE memory, written by the stage:
 *   score:    [0] max key ((ordered f32 bits << 32) | ~global index; compare as unsigned: first argmax)
 *             [1] inverted min key (same rule for the first argmin), [2] sum of the Q.30 weights,
 *             [3..5] (sum, max prefix, min prefix) aggregate of this shard's weights
 *   resample: [0] count key ((offspring count << 32) | ~global index: first most-resampled particle)
 *             [1] first global output slot owned by this shard's sources, [2] one past the last */
int pfmpe_shard_configure(pfmpe_handle* h, int64_t global_offset, int64_t n_global);
/* stage the frame (same constants on every rank; pfmpe_frame_in as for pfmpe_update) */
int pfmpe_shard_begin_frame(pfmpe_handle* h, const pfmpe_frame_in* in);
/* one propagate/project/score pass: `iteration` of the do/while (:535-606) */
int pfmpe_shard_score(pfmpe_handle* h, int iteration, void* d_partial);
/* normalise + stratified resampling (:627-629, :666-682) with the global weight sum, the number of
 * do/while iterations executed and the ordered aggregates of the lower ranks / of all ranks */
int pfmpe_shard_resample(pfmpe_handle* h, int64_t sum_fx, int iterations, const int64_t lower[3], const int64_t all[3],
                         void* d_partial);
/* rows (12 floats: rows 0..2 of the pose) of the output slots [first_slot, first_slot + count);
 * first_owned_slot = partial[1] of pfmpe_shard_resample */
int pfmpe_shard_export(pfmpe_handle* h, int64_t first_owned_slot, int64_t first_slot, int64_t count, void* d_rows);
/* output slots [first_slot, first_slot + count) stay on this rank: they become local particles
 * [local_offset, local_offset + count) of the next frame's set, device to device, no row buffer */
int pfmpe_shard_keep(pfmpe_handle* h, int64_t first_owned_slot, int64_t first_slot, int64_t count, int64_t local_offset);
/* `count` rows received from other ranks become local particles [local_offset, local_offset + count) */
int pfmpe_shard_import(pfmpe_handle* h, const void* d_rows, int64_t local_offset, int64_t count);
/* Peer-memory rebalance (replaces export / all_to_all / keep / import by ONE kernel): every rank
 * publishes the base of its next-frame particle planes (pfmpe_shard_local_ptr within a process,
 * pfmpe_shard_ipc_export -> pfmpe_shard_ipc_open across processes: CUDA IPC, NVLink peer access),
 * pfmpe_shard_set_peers installs the table of all `world` bases (rank order, equal shards) and
 * pfmpe_shard_scatter writes every output slot this rank produced -- `count` slots from
 * first_owned_slot, both from pfmpe_shard_resample's partial -- into the owning rank's planes
 * with local or NVLink peer stores.  Resampling replicates, so a run of slots with the same source
 * inside one destination's range travels as ONE row (stored at the run's first slot) plus a 4-byte
 * leader index per slot (lead[], which lives behind the planes in the same allocation); the next score
 * pass gathers through lead[].  The caller orders the scatter of all ranks before the next
 * pfmpe_shard_score (a barrier on the ranks' streams).  pfmpe_shard_set_peers(h, 0, NULL) removes the
 * table again (row exchange: full rows at every slot). */
int pfmpe_shard_local_ptr(pfmpe_handle* h, void** d_ptr);
int pfmpe_shard_ipc_export(pfmpe_handle* h, unsigned char handle[64]);
int pfmpe_shard_ipc_open(pfmpe_handle* h, const unsigned char handle[64], void** d_ptr);
int pfmpe_shard_set_peers(pfmpe_handle* h, int world, void* const* d_ptrs);
int pfmpe_shard_scatter(pfmpe_handle* h, int64_t first_owned_slot, int64_t count);
/* Peer mailboxes (replace the NCCL all_gather / barrier / broadcast and their device->host copies):
 * every rank creates a mailbox (pfmpe_shard_mail_create returns its device pointer and, for other
 * processes, a CUDA IPC handle to open with pfmpe_shard_ipc_open) and installs the table of all
 * ranks' mailboxes (pfmpe_shard_mail_set_peers, rank order, world <= 32).  From then on
 *   - pfmpe_shard_score / pfmpe_shard_resample deliver their partial from inside the kernel to every
 *     rank with NVLink peer stores; pfmpe_shard_collect(h, -1, table) returns all of them
 *     (table[world][32] int64; word 2 of the resample partial is not exchanged: a rank's end slot is
 *     the next rank's first slot);
 *   - pfmpe_shard_barrier is a stream-ordered barrier across the ranks (no host wait);
 *   - pfmpe_shard_winner_publish (owner only) sends winner pose + correspondences to every rank:
 *     words 0-5 = 12 floats, 6-13 = 16 uint32 pairs, 14 = n; the other ranks call
 *     pfmpe_shard_advance; everybody reads it with pfmpe_shard_collect(h, owner, table).
 * Every rank must take part in every exchange in the same order.  PFMPE_ERR_COMM if a message does
 * not arrive within ~2 s. */
int pfmpe_shard_mail_create(pfmpe_handle* h, int world, int rank, void** d_ptr, unsigned char ipc_handle[64]);
int pfmpe_shard_mail_set_peers(pfmpe_handle* h, void* const* d_ptrs);
int pfmpe_shard_collect(pfmpe_handle* h, int only_rank, int64_t* table);
int pfmpe_shard_advance(pfmpe_handle* h);
int pfmpe_shard_barrier(pfmpe_handle* h);   /* = pfmpe_shard_signal + pfmpe_shard_wait */
int pfmpe_shard_signal(pfmpe_handle* h);    /* the sending half: an empty message to every rank, stream-ordered */
int pfmpe_shard_wait(pfmpe_handle* h);      /* the waiting half: a kernel that returns when every rank's message is in */
int pfmpe_shard_winner_publish(pfmpe_handle* h, int local_index);
/* pose + correspondences (:687-691 / :714-716) of the local particle `local_index` of the last pass;
 * only winner_pose / correspondences / n_correspondences of `out` are meaningful */
int pfmpe_shard_winner(pfmpe_handle* h, int local_index, pfmpe_frame_out* out);

/* One whole PF update of the sharded filter with the host logic in C (pose_estimator.cpp:535-616 do/while + max test,
 * :627-633 sum / accept, :666-688 resampling + winner, taken from the partials of ALL ranks, which the stage kernels
 * deliver into this process's mailbox by NVLink peer stores): score passes until the loop exits, regeneration of the kept
 * iteration, resampling against the global strata, the peer-memory rebalance, a stream-ordered barrier, the winner record
 * from its owner.  `shards` are the handles of this process: one per process in a multi-GPU run (every rank makes the same
 * call with the same inputs and gets the same outputs), or all `world` logical shards of a single-GPU run (they must then
 * share one stream).  Needs pfmpe_shard_configure, pfmpe_shard_set_peers and pfmpe_shard_mail_set_peers on every shard.
 * Shards created with PFMPE_FLAG_MEAN_COV: out->mean_pose / mean_cov of the WHOLE filter (two more exchanges of 12 and 21
 * partial sums through the mailboxes, added in rank order on every rank). */
int pfmpe_shard_frame(pfmpe_handle* const* shards, int n_local, const pfmpe_frame_in* in, pfmpe_frame_out* out);
/* The tracked frame around it, as pfmpe_track_frame_host does for an unsharded filter: predictPose (host, FP64), the
 * sharded PF update, optimiseAndUpdatePose + jump check (host, FP64); pfmpe_track_init on every shard first. */
int pfmpe_shard_track_frame_host(pfmpe_handle* const* shards, int n_local, double time_to_predict, const double* image_points,
                                 int32_t n_image_points, uint32_t seed, pfmpe_track_result* out);
/* score passes pfmpe_shard_frame has launched on this shard so far */
int pfmpe_shard_passes(const pfmpe_handle* h, uint64_t* n);

/* ---- full-size parity aid ------------------------------------------------------------------ */
/* Position-keyed 64-bit checksums of one tracker's device state, computed on the device: [0] PoseParticle (propagated
 * set of the kept iteration), [1] its raw weights, [2] newPoseEstimation (the resampled set).  Every word is mixed with
 * its GLOBAL particle / slot number and the results are summed mod 2^64, so the checksums of the shards of a sharded
 * filter add up to those of the unsharded filter exactly when the states are identical (no host copy of 2^26 poses). */
int pfmpe_state_checksum(pfmpe_handle* h, int tracker, uint64_t out[3]);

/* ---- streams / timing ---------------------------------------------------------------------- */
int pfmpe_set_stream(pfmpe_handle* h, void* cuda_stream); /* use an external cudaStream_t (e.g. torch's) */
int pfmpe_sync(pfmpe_handle* h);
/* device time of the most recent pfmpe_track_run / pfmpe_track_replay launch (CUDA events on the handle's
 * stream); the per-frame calls (pfmpe_update*) are the latency path and record no events */
int pfmpe_last_launch_ms(pfmpe_handle* h, float* ms);
/* number of kernels this handle has launched so far */
int pfmpe_launch_count(const pfmpe_handle* h, uint64_t* n);
/* launch geometry chosen for the hot-path kernel: grid.x, grid.y, block, cooperative? */
int pfmpe_launch_info(const pfmpe_handle* h, int32_t info[4]);

/* Profiling aid: enable/disable in-kernel %globaltimer stamps (CTA 0, thread 0) and read the stamps of the
 * most recent frame: [0] frame start, [1] blobs staged, [2] first score pass done (this CTA), [3] after the
 * grid reduction, [4] scan start, [5] scan done (this CTA), [6] after the count reduction, [7] result record
 * written, [8] tail done (incl. device Gauss-Newton in track mode), [9] after the end-of-frame barrier. */
int pfmpe_debug_timeline(pfmpe_handle* h, int enable, uint64_t stamps_ns[16]);

#ifdef __cplusplus
}
#endif
#endif /* PFMPE_H_ */
