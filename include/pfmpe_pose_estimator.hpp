/*
 * pfmpe_pose_estimator.hpp -- C++ host-side mirror of the reference's PoseEstimator interface for the PF path
 * (pf_mpe_lib/include/pf_mpe_lib/pose_estimator.h), header-only, on top of the C ABI (pfmpe.h, libpfmpe.so).
 *
 * Same member names, argument meaning and error behaviour as the reference: everything by value,
 * no exceptions, no return codes -- failure is `false` in the vector estimateBodyPose returns and a float code in
 * getPublisherData().Flag_Fail[object].  The types are dependency-free stand-ins with the reference's memory
 * layouts (Eigen / OpenCV / ROS are not needed): Matrix4d = 16 column-major doubles (Eigen::Matrix4d), the camera
 * matrix = 9 row-major doubles (cv::Mat at<double>(i,j)).  Differences, all outside the hot path:
 *   - estimateBodyPose takes the detected, undistorted LED positions of each object (what LEDDetector::findLeds
 *     + setImagePoints hand to the PF branch at :449-469) instead of the camera image: detection is out of scope;
 *   - after initialise() the reference seeds the filter from its P3P estimates; here setInitialPoses() provides
 *     the two poses the tracker starts from (the state :180-191 leaves behind) and correspondenceHistogram()
 *     exposes the device-side vote of initialise() (:1524-1717) for a caller that runs the rest of initialise();
 *   - setSeed() replaces std::random_device (:476-477) when a reproducible stream is wanted.
 * There is no CPU fallback: without a CUDA device every estimateBodyPose() fails with Flag_Fail = 13.
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

  struct PubData_struct {                            // :158-167
    float Flag_Fail[2];
    int numIter;
    double highestProb;
    PubData_struct() : numIter(0), highestProb(0) { Flag_Fail[0] = Flag_Fail[1] = 0.f; }
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
  // setMarkerPositions (:56-62): the first argument is the marker set of object 0, the vector holds every object's
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
    return check(pfmpe_track_init(h_, objectNumber, pose0.data(), t0, pose1.data(), t1));
  }

  // ---- the per-frame call (:69-896, PF branch) ------------------------------------------------------------------
  // detected_led_positions[object]: undistorted pixel positions of the blobs found in that object's ROI.
  std::vector<bool> estimateBodyPose(const std::vector<List2DPoints>& detected_led_positions, double time_to_predict)
  {
    std::vector<bool> found(object_points_Vec.size(), false);
    if (!ensure()) {
      for (size_t o = 0; o < found.size() && o < 2; ++o) PubData.Flag_Fail[o] = 13.f;   // no device / bad configuration
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
    const int rc = T == 1 ? pfmpe_track_frame_host(h_, 0, time_to_predict, pts.data(), counts[0], seeds[0], &results_[0])
                          : pfmpe_track_frame_host_batch(h_, time_to_predict, pts.data(), counts.data(), seeds.data(), results_.data());
    if (!check(rc)) {
      for (int o = 0; o < T && o < 2; ++o) PubData.Flag_Fail[o] = 13.f;
      return found;
    }
    for (int o = 0; o < T; ++o) {
      found[o] = results_[o].status == 1 && results_[o].flag_fail < 4.f;
      if (o < 2) PubData.Flag_Fail[o] = results_[o].flag_fail;
    }
    PubData.numIter = results_[0].gn_iterations;
    PubData.highestProb = results_[0].highest_prob;
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
  PubData_struct getPublisherData() { return PubData; }                                          // :1798

  // ROI corner points for LEDDetector::determineROI (predictMarkerPositionsInImage :1036-1052 + :221-244)
  bool predictROI(int objectNumber, const Matrix4d& camMoveInv, const Matrix4d& predictionMatrix, const Matrix4d& predicted_pose,
                  double roi[4])
  {
    return ensure() && check(pfmpe_predict_roi(h_, objectNumber, camMoveInv.data(), predictionMatrix.data(), predicted_pose.data(), roi));
  }

  // hist_corr of initialise() (:1524-1717): rows = detections, columns = LEDs (row-major)
  bool correspondenceHistogram(int objectNumber, const List2DPoints& image_points, std::vector<uint32_t>& hist_corr)
  {
    if (!ensure() || image_points.empty()) return false;
    sync_params();
    std::vector<double> pts(2 * image_points.size());
    for (size_t i = 0; i < image_points.size(); ++i) { pts[2 * i] = image_points[i][0]; pts[2 * i + 1] = image_points[i][1]; }
    hist_corr.assign(image_points.size() * (size_t)built_leds_, 0u);
    return check(pfmpe_p3p_histogram(h_, objectNumber, pts.data(), (int32_t)image_points.size(), hist_corr.data(), NULL));
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
  // (re)create the handle when the shape changed: N_Particle, number of objects, LEDs per object, device
  bool ensure()
  {
    const int T = (int)object_points_Vec.size(), L = T ? (int)object_points_Vec[0].size() : 0;
    if (h_ && !dirty_ && built_particles_ == N_Particle && built_objects_ == T && built_leds_ == L) return true;
    if (h_) { pfmpe_destroy(h_); h_ = NULL; }
    if (!bUseParticleFilter || T < 1 || L < 1 || L > PFMPE_MAX_LED || N_Particle < 2) { err_ = "unsupported configuration"; return false; }
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
  bool have_seed_;
  uint32_t seed_, frame_ = 0;
  std::random_device rd_;
  std::string err_;
};

}  // namespace pfmpe
#endif /* PFMPE_POSE_ESTIMATOR_HPP_ */
