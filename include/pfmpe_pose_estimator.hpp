/*
 * pfmpe_pose_estimator.hpp -- C++ host-side mirror of the reference's PoseEstimator interface for the PF path
 * (pf_mpe_lib/include/pf_mpe_lib/pose_estimator.h), header-only, on top of the C ABI (pfmpe.h, libpfmpe.so).
 *
 * Same member names, argument meaning and error behaviour as the reference: everything by value,
 * no exceptions, no return codes -- failure is `false` in the vector estimateBodyPose returns and a float code in
 * getPublisherData().Flag_Fail[object].  The types are dependency-free stand-ins with the reference's memory
 * layouts (Eigen / OpenCV / ROS are not needed): Matrix4d = 16 column-major doubles (Eigen::Matrix4d), the camera
 * matrix = 9 row-major doubles (cv::Mat at<double>(i,j)); PubData_struct has the reference's fields in the
 * reference's order (pose_estimator.h:158-167), so a node compiled against this header reads the same record.
 * Differences, all outside the hot path:
 *   - estimateBodyPose takes the detected, undistorted LED positions of each object (what LEDDetector::findLeds
 *     + setImagePoints hand to the PF branch at :449-469) instead of the camera image: detection is out of scope;
 *   - the initialisation branch (:149-187) keeps the reference's own host code (initialise(), correspondencesFromHistogram,
 *     the accept logic of checkCorrespondences, computeTransformation: SURVEY 2 #8, out of scope).  Only its two
 *     data-parallel loops have device stand-ins here -- correspondenceHistogram() for the vote (:1524-1717) and
 *     checkCorrespondenceSets() for the per-combination bodies (:1340-1445) -- and seedInitialisedObject() hands what the
 *     reference's initialise() + optimiseAndUpdatePose leave behind (:160-186) to the filter; setInitialPoses() provides
 *     a known start instead;
 *   - setSeed() replaces std::random_device (:476-477) when a reproducible stream is wanted;
 *   - the library keeps previous / current time per tracked object (the reference shares current_time_ / previous_time_
 *     between objects, so that objects >= 1 of a frame see predicted_time_ == current_time_ and an identity
 *     predictionMatrix, :995-1010 after :2016-2020): a deliberate deviation, see DESIGN.md section 2.
 * There is no CPU fallback: without a CUDA device every call fails and Flag_Fail = kFlagFailNoDevice (100; the
 * reference's own codes are -1 .. 16, pf_mpe/include/pf_mpe/monocular_pose_estimator.h:119-140).
 */
#ifndef PFMPE_POSE_ESTIMATOR_HPP_
#define PFMPE_POSE_ESTIMATOR_HPP_

#include <array>
#include <cstdint>
#include <random>
#include <string>
#include <utility>
#include <vector>

#include "pfmpe.h"

namespace pfmpe {

typedef std::array<double, 16> Matrix4d;                       // column-major, Eigen::Matrix4d layout
typedef std::array<double, 36> Matrix6d;                       // column-major
typedef std::array<double, 2> Vector2d;
typedef std::array<double, 4> Vector4d;                        // homogeneous (x, y, z, 1)
typedef std::vector<Vector2d> List2DPoints;                    // pose_estimator.h: List2DPoints
typedef std::vector<Vector4d> List4DPoints;                    // pose_estimator.h: List4DPoints
typedef std::vector<std::pair<unsigned, unsigned> > VectorXuPairs;   // rows (LED, detection), 1-based

static const float kFlagFailNoDevice = 100.f;                  // not a reference code: no CUDA device / C-ABI failure

class PoseEstimator {
 public:
  // ---- public members of the reference (pose_estimator.h:121-155) that the PF path reads -----------------------
  std::array<double, 9> camera_matrix_K_;            // :121, row-major 3x3
  std::vector<double> camera_distortion_coeffs_;     // :122 (used by the plumb_bob option only)
  bool bUseParticleFilter;                           // :138 -- must stay true: this IS the particle filter
  int N_Particle;                                    // :139
  double maxAngularNoise, minAngularNoise;           // :140-141
  double maxTransitionNoise, minTransitionNoise;     // :142-143
  double back_projection_pixel_tolerance_PF;         // :147
  bool bMarkerDowngrade[5];                          // :149
  bool bUseCamPos;                                   // :127 (false: camMoveInv = I)

  struct PubData_struct {                            // :158-167, same fields, order and types
    float bPred;                                     // 1: ROI prediction was used, 0: brute force (:165, :2067)
    float numIter;                                   // Gauss-Newton iterations of optimisePose (:1883)
    float numDetLED;                                 // detected_led_positions.size() (:164)
    float EstProb[4];                                // setP3PEstProb (:2319, "no more used" in the reference)
    float Flag_Fail[2];
    PubData_struct() : bPred(0.f), numIter(0.f), numDetLED(0.f) { EstProb[0] = EstProb[1] = EstProb[2] = EstProb[3] = 0.f; Flag_Fail[0] = Flag_Fail[1] = 0.f; }
  };

  PoseEstimator()
      : bUseParticleFilter(true), N_Particle(100), maxAngularNoise(0.015), minAngularNoise(-0.015),
        maxTransitionNoise(0.035), minTransitionNoise(-0.035), back_projection_pixel_tolerance_PF(4.0), bUseCamPos(false),
        back_projection_pixel_tolerance_(5.0), h_(NULL), built_particles_(0), built_objects_(0), built_leds_(0),
        device_(0), have_seed_(false), seed_(0)
  {
    camera_matrix_K_.fill(0.0);
    camera_matrix_K_[0] = camera_matrix_K_[4] = camera_matrix_K_[8] = 1.0;
    for (int i = 0; i < 5; ++i) bMarkerDowngrade[i] = false;
  }
  ~PoseEstimator() { if (h_) pfmpe_destroy(h_); }

  // ---- configuration (same names as the reference) -------------------------------------------------------------
  // setMarkerPositions (:56-62): the first argument is the marker set of object 0, the vector holds every object's.
  // All objects must carry the same number of LEDs (one kernel instantiation per handle); otherwise every later call
  // fails cleanly with lastError() = "objects with different LED counts".
  void setMarkerPositions(List4DPoints positions_of_markers_on_object, std::vector<List4DPoints> positions_of_markers_on_object_vector)
  {
    object_points_ = positions_of_markers_on_object;
    object_points_Vec = positions_of_markers_on_object_vector;
    if (object_points_Vec.empty()) object_points_Vec.push_back(object_points_);
    dirty_ = true;
  }
  List4DPoints getMarkerPositions() { return object_points_; }                                  // :64
  void setBackProjectionPixelTolerance(double tolerance) { back_projection_pixel_tolerance_ = tolerance; }   // :945
  double getBackProjectionPixelTolerance() { return back_projection_pixel_tolerance_; }                       // :950
  void setDevice(int cuda_ordinal) { device_ = cuda_ordinal; dirty_ = true; }
  void setSeed(uint32_t seed) { have_seed_ = true; seed_ = seed; }     // per-frame seed = seed + frame counter

  // The state initialise() + the first tracked frame leave behind (:180-191): previous / current pose and times,
  // it_since_initialized_ = 2, every particle = pose1.
  bool setInitialPoses(int objectNumber, const Matrix4d& pose0, double t0, const Matrix4d& pose1, double t1)
  {
    if (!ensure()) return false;
    if (objectNumber < 0 || objectNumber >= built_objects_) return false;
    results_[objectNumber] = pfmpe_track_result();
    for (int i = 0; i < 16; ++i) results_[objectNumber].pose[i] = pose1[i];
    if (!check(pfmpe_track_init(h_, objectNumber, pose0.data(), t0, pose1.data(), t1))) return false;
    previous_time_ = t0; current_time_ = t1;
    return true;
  }

  // ---- the per-frame call (:69-896, PF branch) ------------------------------------------------------------------
  // detected_led_positions[object]: undistorted pixel positions of the blobs found in that object's ROI.
  std::vector<bool> estimateBodyPose(const std::vector<List2DPoints>& detected_led_positions, double time_to_predict)
  {
    std::vector<bool> found(object_points_Vec.size(), false);
    if (!ensure()) {
      for (size_t o = 0; o < found.size() && o < 2; ++o) PubData.Flag_Fail[o] = kFlagFailNoDevice;
      return found;
    }
    sync_params();
    const int T = built_objects_;
    std::vector<double> pts((size_t)T * max_blobs_ * 2, 0.0);
    std::vector<int32_t> counts(T, 0);
    std::vector<uint32_t> seeds(T, 0);
    for (int o = 0; o < T; ++o) {
      const List2DPoints empty;
      const List2DPoints& d = o < (int)detected_led_positions.size() ? detected_led_positions[o] : empty;
      const int n = (int)d.size() < max_blobs_ ? (int)d.size() : max_blobs_;
      counts[o] = n;
      for (int i = 0; i < n; ++i) { pts[((size_t)o * max_blobs_ + i) * 2] = d[i][0]; pts[((size_t)o * max_blobs_ + i) * 2 + 1] = d[i][1]; }
      seeds[o] = have_seed_ ? seed_ + frame_ + 100003u * (uint32_t)o : rd_();          // :476-477
    }
    ++frame_;
    image_points_ = detected_led_positions.empty() ? List2DPoints() : detected_led_positions[0];
    PubData.numDetLED = (float)image_points_.size();
    PubData.bPred = 1.f;                                                               // tracking branch: the prediction was used
    const int rc = T == 1 ? pfmpe_track_frame_host(h_, 0, time_to_predict, pts.data(), counts[0], seeds[0], &results_[0])
                          : pfmpe_track_frame_host_batch(h_, time_to_predict, pts.data(), counts.data(), seeds.data(), results_.data());
    if (!check(rc)) {
      for (int o = 0; o < T && o < 2; ++o) PubData.Flag_Fail[o] = kFlagFailNoDevice;
      return found;
    }
    bool any_updated = false;
    for (int o = 0; o < T; ++o) {
      found[o] = results_[o].status == 1 && results_[o].flag_fail < 4.f;
      any_updated = any_updated || results_[o].most_likely_pose_idx >= 0;              // resampled: optimiseAndUpdatePose ran (:690)
      if (o < 2) PubData.Flag_Fail[o] = results_[o].flag_fail;
    }
    if (any_updated) advance_time(time_to_predict);                                    // updatePose (:2016-2020)
    PubData.numIter = (float)results_[0].gn_iterations;
    highest_prob_ = results_[0].highest_prob;
    return found;
  }

  // ---- results (same names as the reference) --------------------------------------------------------------------
  Matrix4d getPredictedPose(int objectNumber)                                                    // :905
  {
    Matrix4d m; m.fill(0.0);
    if (objectNumber >= 0 && objectNumber < (int)results_.size()) for (int i = 0; i < 16; ++i) m[i] = results_[objectNumber].pose[i];
    return m;
  }
  Matrix6d getPoseCovariance(int objectNumber)                                                   // :911
  {
    Matrix6d m; m.fill(0.0);
    if (objectNumber >= 0 && objectNumber < (int)results_.size()) for (int i = 0; i < 36; ++i) m[i] = results_[objectNumber].cov[i];
    return m;
  }
  std::vector<Matrix4d> getPoseParticles(int objectNumber) { return particles(objectNumber, 0); }        // :917
  std::vector<Matrix4d> getResampledParticles(int objectNumber) { return particles(objectNumber, 1); }   // :923
  std::vector<double> getParticleWeights(int objectNumber)                                       // probPart_Vec (:733)
  {
    std::vector<double> w;
    if (!h_ || objectNumber < 0 || objectNumber >= built_objects_) return w;
    w.resize(built_particles_);
    if (!check(pfmpe_get_weights(h_, objectNumber, 1, w.data()))) w.clear();
    return w;
  }
  VectorXuPairs getCorrespondences()                                                             // :1067
  {
    VectorXuPairs c;
    if (results_.empty()) return c;
    for (int i = 0; i < results_[0].n_correspondences; ++i)
      c.push_back(std::make_pair(results_[0].correspondences[2 * i], results_[0].correspondences[2 * i + 1]));
    return c;
  }
  List2DPoints getImagePoints() { return image_points_; }                                        // :1012
  void setImagePoints(List2DPoints points) { image_points_ = points; }                           // :929
  PubData_struct getPublisherData() { return PubData; }                                          // :1798
  double getHighestProbability() const { return highest_prob_; }                                 // highestProb (:519, :610) of object 0

  // ROI corner points for LEDDetector::determineROI (predictMarkerPositionsInImage :1036-1052 + :221-244)
  bool predictROI(int objectNumber, const Matrix4d& camMoveInv, const Matrix4d& predictionMatrix, const Matrix4d& predicted_pose,
                  double roi[4])
  {
    return ensure() && check(pfmpe_predict_roi(h_, objectNumber, camMoveInv.data(), predictionMatrix.data(), predicted_pose.data(), roi));
  }

  // ---- device stand-ins for the two data-parallel loops of the brute-force initialisation -----------------------
  // hist_corr of initialise() (:1524-1717): rows = detections, columns = LEDs (row-major); the reference feeds it to its
  // own correspondencesFromHistogram (:1134-1288) unchanged.
  bool correspondenceHistogram(int objectNumber, const List2DPoints& image_points, std::vector<uint32_t>& hist_corr)
  {
    if (!ensure() || image_points.empty() || objectNumber < 0 || objectNumber >= built_objects_) return false;
    sync_params();
    std::vector<double> pts;
    flatten(image_points, pts);
    hist_corr.assign(image_points.size() * (size_t)built_leds_, 0u);
    return check(pfmpe_p3p_histogram(h_, objectNumber, pts.data(), (int32_t)image_points.size(), hist_corr.data(), NULL));
  }

  // The per-combination bodies of checkCorrespondences (:1340-1445) for a list of candidate correspondence sets (the
  // entries correspondencesFromHistogram returns, :1733): results[s * stride + c] is the record of combination c
  // (lexicographic = Combinations::combinationsNoReplacement order) of set s; stride = C(n_leds, 3).  The reference's
  // accept logic (:1427-1488) consumes the records.
  bool checkCorrespondenceSets(int objectNumber, const List2DPoints& image_points, const std::vector<VectorXuPairs>& sets,
                               std::vector<pfmpe_p3p_check_result>& results, int& stride)
  {
    if (!ensure() || objectNumber < 0 || objectNumber >= built_objects_) return false;
    sync_params();
    const int L = built_leds_;
    stride = L * (L - 1) * (L - 2) / 6;
    std::vector<double> pts;
    flatten(image_points, pts);
    const int chunk_sets = 16384;
    results.assign(sets.size() * (size_t)stride, pfmpe_p3p_check_result());
    std::vector<uint32_t> rows;
    std::vector<int32_t> counts;
    for (size_t s0 = 0; s0 < sets.size(); s0 += chunk_sets) {
      const int m = (int)(sets.size() - s0 < (size_t)chunk_sets ? sets.size() - s0 : (size_t)chunk_sets);
      rows.assign((size_t)m * 2 * PFMPE_MAX_LED, 0u);
      counts.assign(m, 0);
      for (int s = 0; s < m; ++s) {
        const VectorXuPairs& c = sets[s0 + s];
        counts[s] = (int32_t)(c.size() < (size_t)PFMPE_MAX_LED ? c.size() : (size_t)PFMPE_MAX_LED);
        for (int i = 0; i < counts[s]; ++i) { rows[((size_t)s * PFMPE_MAX_LED + i) * 2] = c[i].first; rows[((size_t)s * PFMPE_MAX_LED + i) * 2 + 1] = c[i].second; }
      }
      if (!check(pfmpe_p3p_check(h_, objectNumber, pts.data(), (int32_t)image_points.size(), rows.data(), counts.data(), m, stride,
                                 certainty_threshold_, &results[s0 * (size_t)stride])))
        return false;
    }
    return true;
  }
  void setCertaintyThreshold(double threshold) { certainty_threshold_ = threshold; }                    // :965
  double getCertaintyThreshold() { return certainty_threshold_; }                                        // :970

  // optimisePose (:1805-2009) on the given image points and correspondences; pose in/out
  bool optimisePose(int objectNumber, const List2DPoints& image_points, const VectorXuPairs& correspondences, Matrix4d& pose,
                    Matrix6d& cov, int& iterations)
  {
    if (objectNumber < 0 || objectNumber >= (int)object_points_Vec.size()) return false;
    const List4DPoints& obj = object_points_Vec[objectNumber];
    std::vector<double> xyz(3 * obj.size()), pts;
    for (size_t j = 0; j < obj.size(); ++j) for (int c = 0; c < 3; ++c) xyz[3 * j + c] = obj[j][c];
    flatten(image_points, pts);
    std::vector<uint32_t> corr(2 * correspondences.size());
    for (size_t i = 0; i < correspondences.size(); ++i) { corr[2 * i] = correspondences[i].first; corr[2 * i + 1] = correspondences[i].second; }
    return check(pfmpe_optimise_pose(camera_matrix_K_.data(), xyz.data(), (int)obj.size(), pts.data(), (int)image_points.size(), corr.data(),
                                     (int)correspondences.size(), pose.data(), cov.data(), &iterations));
  }

  // What a successful initialise() leaves behind in the reference (:160-186), handed to the filter: PoseParticle_Vec =
  // newPoseEstimation_Vec = the P3P seeds, previous_pose_ = the unoptimised pose (current_pose_ = predicted_pose_, :160),
  // current_pose_ = the optimised one, times by updatePose's rule (:2016-2020), it_since_initialized_ = 1, Flag_Fail = 0.
  // The next estimateBodyPose() runs the particle filter from there.
  bool seedInitialisedObject(int objectNumber, const std::vector<Matrix4d>& PoseParticle, const Matrix4d& unoptimised_pose,
                             const Matrix4d& optimised_pose, const Matrix6d& pose_covariance, int numIter, double time_to_predict)
  {
    if (!ensure() || objectNumber < 0 || objectNumber >= built_objects_ || (int)PoseParticle.size() != built_particles_) {
      setFlag(objectNumber, kFlagFailNoDevice);
      return false;
    }
    std::vector<double> flat((size_t)built_particles_ * 16);
    for (int n = 0; n < built_particles_; ++n) for (int i = 0; i < 16; ++i) flat[(size_t)n * 16 + i] = PoseParticle[(size_t)n][i];
    advance_time(time_to_predict);
    if (!check(pfmpe_set_particles(h_, objectNumber, flat.data(), built_particles_)) ||
        !check(pfmpe_track_set_state(h_, objectNumber, unoptimised_pose.data(), previous_time_, optimised_pose.data(), current_time_, 1u))) {
      setFlag(objectNumber, kFlagFailNoDevice);
      return false;
    }
    results_[objectNumber] = pfmpe_track_result();
    for (int i = 0; i < 16; ++i) results_[objectNumber].pose[i] = optimised_pose[i];
    for (int i = 0; i < 36; ++i) results_[objectNumber].cov[i] = pose_covariance[i];
    PubData.numIter = (float)numIter;
    PubData.bPred = 0.f;                                                                   // brute force approach was used (:165)
    setFlag(objectNumber, 0.f);                                                            // :195
    return true;
  }

  bool ok() const { return h_ != NULL; }
  const std::string& lastError() const { return err_; }

 private:
  bool check(int rc)
  {
    if (rc == PFMPE_OK) return true;
    const char* d = h_ ? pfmpe_last_error(h_) : "";
    err_ = std::string(pfmpe_strerror(rc)) + (d && *d ? std::string(": ") + d : std::string());
    return false;
  }
  // updatePose's time bookkeeping (:2016-2020); members shared between the objects like the reference's
  void advance_time(double predicted_time)
  {
    if (predicted_time - current_time_ > 0.001 || predicted_time < current_time_) { previous_time_ = current_time_; current_time_ = predicted_time; }
  }
  static void flatten(const List2DPoints& p, std::vector<double>& out)
  {
    out.resize(2 * p.size());
    for (size_t i = 0; i < p.size(); ++i) { out[2 * i] = p[i][0]; out[2 * i + 1] = p[i][1]; }
  }
  // (re)create the handle when the shape changed: N_Particle, number of objects, LEDs per object, device
  bool ensure()
  {
    const int T = (int)object_points_Vec.size(), L = T ? (int)object_points_Vec[0].size() : 0;
    if (h_ && !dirty_ && built_particles_ == N_Particle && built_objects_ == T && built_leds_ == L) return true;
    if (h_) { pfmpe_destroy(h_); h_ = NULL; }
    if (!bUseParticleFilter || T < 1 || L < 1 || L > PFMPE_MAX_LED || N_Particle < 2) { err_ = "unsupported configuration"; return false; }
    for (int o = 1; o < T; ++o)
      if ((int)object_points_Vec[o].size() != L) { err_ = "unsupported configuration: objects with different LED counts"; return false; }
    pfmpe_config cfg = pfmpe_config();
    cfg.n_particles = N_Particle; cfg.n_trackers = T; cfg.n_leds = L; cfg.max_blobs = max_blobs_; cfg.device = device_;
    const int rc = pfmpe_create(&cfg, &h_);
    if (rc != PFMPE_OK) { h_ = NULL; err_ = std::string(pfmpe_strerror(rc)) + ": " + pfmpe_last_error(NULL); return false; }
    for (int o = 0; o < T; ++o) {
      std::vector<double> xyz(3 * (size_t)L);
      for (int j = 0; j < L; ++j) for (int c = 0; c < 3; ++c) xyz[3 * j + c] = object_points_Vec[o][j][c];
      if (!check(pfmpe_set_markers(h_, o, xyz.data(), L))) { pfmpe_destroy(h_); h_ = NULL; return false; }
    }
    built_particles_ = N_Particle; built_objects_ = T; built_leds_ = L; dirty_ = false;
    results_.assign(T, pfmpe_track_result());
    frame_ = 0;
    sync_params();
    return true;
  }
  // the public fields may change at any time (dynamic_reconfigure, monocular_pose_estimator.cpp:480-527)
  void sync_params()
  {
    pfmpe_set_camera(h_, camera_matrix_K_.data(), camera_distortion_coeffs_.size() >= 5 ? camera_distortion_coeffs_.data() : NULL);
    pfmpe_pf_params p = pfmpe_pf_params();
    p.min_angular_noise = minAngularNoise; p.max_angular_noise = maxAngularNoise;
    p.min_transition_noise = minTransitionNoise; p.max_transition_noise = maxTransitionNoise;
    p.back_projection_pixel_tolerance = back_projection_pixel_tolerance_;
    p.back_projection_pixel_tolerance_PF = back_projection_pixel_tolerance_PF;
    for (int i = 0; i < 5; ++i) p.marker_downgrade[i] = bMarkerDowngrade[i] ? 1 : 0;
    pfmpe_set_params(h_, &p);
  }
  void setFlag(int objectNumber, float f) { if (objectNumber >= 0 && objectNumber < 2) PubData.Flag_Fail[objectNumber] = f; }
  std::vector<Matrix4d> particles(int objectNumber, int which)
  {
    std::vector<Matrix4d> out;
    if (!h_ || objectNumber < 0 || objectNumber >= built_objects_) return out;
    std::vector<double> buf((size_t)built_particles_ * 16);
    if (!check(pfmpe_get_particles(h_, objectNumber, which, buf.data()))) return out;
    out.resize(built_particles_);
    for (int n = 0; n < built_particles_; ++n) for (int i = 0; i < 16; ++i) out[n][i] = buf[(size_t)n * 16 + i];
    return out;
  }

  static const int max_blobs_ = 256;                 // capacity for detections per object and frame
  List4DPoints object_points_;
  std::vector<List4DPoints> object_points_Vec;
  List2DPoints image_points_;
  double back_projection_pixel_tolerance_;
  PubData_struct PubData;
  pfmpe_handle* h_;
  bool dirty_ = true;
  int built_particles_, built_objects_, built_leds_, device_;
  std::vector<pfmpe_track_result> results_;
  double certainty_threshold_ = 1.0;                 // cfg default (PFMonocularPoseEstimator.cfg:20)
  double previous_time_ = 0.0, current_time_ = 0.0;  // :2016-2020
  double highest_prob_ = 0.0;
  bool have_seed_;
  uint32_t seed_, frame_ = 0;
  std::random_device rd_;
  std::string err_;
};

}  // namespace pfmpe
#endif /* PFMPE_POSE_ESTIMATOR_HPP_ */
