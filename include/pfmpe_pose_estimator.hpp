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
 *   - the initialisation branch of estimateBodyPose (:149-187) is its own call, initialiseBodyPose(): device vote
 *     (:1524-1717) and checkCorrespondences bodies (:1340-1445), host correspondencesFromHistogram / accept logic /
 *     computeTransformation / particle seeding as in the reference; setInitialPoses() instead provides a known start;
 *   - setSeed() replaces std::random_device (:476-477) when a reproducible stream is wanted.
 * There is no CPU fallback: without a CUDA device every estimateBodyPose() fails with Flag_Fail = 13.
 */
#ifndef PFMPE_POSE_ESTIMATOR_HPP_
#define PFMPE_POSE_ESTIMATOR_HPP_

#include <algorithm>
#include <array>
#include <cmath>
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

  // ---- brute-force initialisation (:1503-1786 + the branch of estimateBodyPose that calls it, :149-187) ---------
  void setCertaintyThreshold(double threshold) { certainty_threshold_ = threshold; }                    // :965
  double getCertaintyThreshold() { return certainty_threshold_; }                                        // :970
  void setValidCorrespondenceThreshold(double threshold) { valid_correspondence_threshold_ = threshold; }   // :975
  double getValidCorrespondenceThreshold() { return valid_correspondence_threshold_; }                   // :980

  // correspondencesFromHistogram (:1134-1288): candidate correspondence sets, most probable first.
  // histogram: numRows x numCols (detections x LEDs), row-major.
  static std::vector<VectorXuPairs> correspondencesFromHistogram(const std::vector<uint32_t>& histogram, unsigned numRows,
                                                                 unsigned numCols, bool bInitialisation)
  {
    std::vector<VectorXuPairs> correspondencesVec;
    const double prob_threshold = (1.3 * 1.0) / (numRows * numCols);
    std::vector<double> hp(histogram.begin(), histogram.end());
    std::vector<unsigned> rowSum(numRows, 0u);
    for (unsigned r = 0; r < numRows; ++r) for (unsigned c = 0; c < numCols; ++c) rowSum[r] += histogram[r * numCols + c];
    for (unsigned c = 0; c < numCols; ++c) {
      unsigned colSum = 0;
      for (unsigned r = 0; r < numRows; ++r) colSum += histogram[r * numCols + c];
      if (colSum == 0) continue;                                    // no detection close enough for this LED
      for (unsigned r = 0; r < numRows; ++r) {
        double& v = hp[r * numCols + c];
        v = std::max(0.0, std::pow(v, 2) / (colSum * rowSum[r]));  // unsigned product, as in the reference (:1155)
        if (v < prob_threshold) v = 0;
      }
    }
    std::vector<std::vector<double> > u_prob(numCols);
    std::vector<std::vector<int> > u_num(numCols);
    for (unsigned a = 0; a < numCols; ++a)
      for (unsigned b = 0; b < numRows; ++b)
        if (hp[b * numCols + a] != 0) { u_prob[a].push_back(hp[b * numCols + a]); u_num[a].push_back((int)b + 1); }
    long long N = 1;
    std::vector<int> N_v(numCols);
    for (unsigned k = 0; k < numCols; ++k) {
      N *= std::max<long long>(1, (long long)u_prob[k].size());
      N_v[k] = (int)u_num[k].size();
      if (N > (1ll << 22)) return correspondencesVec;               // the reference would enumerate them all; refuse instead
    }
    std::vector<std::vector<int> > v_comb((size_t)N);
    std::vector<double> v_prob_comb((size_t)N);
    for (long long i = 0; i < N; ++i) {
      double prob = 1;
      long long n = 1;
      std::vector<int> comb;
      for (int idxLED = (int)numCols - 1; idxLED > -1; --idxLED) {
        if (N_v[idxLED] > 0) {
          const int idxDet = (int)((i / n) % N_v[idxLED]);
          prob *= u_prob[idxLED][idxDet];
          comb.push_back(u_num[idxLED][idxDet]);
          n *= std::max(1, N_v[idxLED]);
        } else comb.push_back(0);
      }
      std::reverse(comb.begin(), comb.end());
      v_prob_comb[(size_t)i] = prob;
      v_comb[(size_t)i] = comb;
    }
    double sum = 0;
    for (size_t i = 0; i < v_prob_comb.size(); ++i) sum += v_prob_comb[i];
    for (size_t i = 0; i < v_prob_comb.size(); ++i) v_prob_comb[i] = v_prob_comb[i] / sum;
    // the reference picks max_element N times (first maximum wins, O(N^2)); a stable sort by descending probability
    // visits the candidates in exactly that order
    std::vector<size_t> order(v_prob_comb.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return v_prob_comb[a] > v_prob_comb[b]; });
    for (size_t bb = 0; bb < order.size(); ++bb) {
      const std::vector<int>& det = v_comb[order[bb]];
      if (bInitialisation && checkAmbiguity(det)) continue;         // occlusions are only allowed when re-initialising
      VectorXuPairs pairs;
      for (unsigned led = 0; led < numCols; ++led)
        if (det[led] != 0) pairs.push_back(std::make_pair(led + 1, (unsigned)det[led]));
      correspondencesVec.push_back(pairs);
    }
    return correspondencesVec;
  }

  // checkAmbiguity (:2447-2458): two LEDs claim the same detection (the reference also reads one element past the
  // end of the vector, which is not reproduced)
  static bool checkAmbiguity(const std::vector<int>& correspondingDetections)
  {
    for (size_t i = 0; i + 1 < correspondingDetections.size(); ++i)
      for (size_t j = correspondingDetections.size() - 1; j > i; --j)
        if (correspondingDetections[i] == correspondingDetections[j]) return true;
    return false;
  }

  // computeTransformation (:2139-2161): rigid transform between two 3 x n point sets (columns), R = V U^T of the SVD
  // of the cross-covariance (Jacobi on H^T H; a vanishing third singular value is completed to a proper rotation)
  static Matrix4d computeTransformation(const std::vector<std::array<double, 3> >& object_points,
                                        const std::vector<std::array<double, 3> >& reprojected_points)
  {
    const size_t n = object_points.size();
    double mo[3] = {0, 0, 0}, mr[3] = {0, 0, 0};
    for (size_t i = 0; i < n; ++i) for (int k = 0; k < 3; ++k) { mo[k] += object_points[i][k]; mr[k] += reprojected_points[i][k]; }
    for (int k = 0; k < 3; ++k) { mo[k] /= (double)n; mr[k] /= (double)n; }
    double H[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
    for (size_t i = 0; i < n; ++i)
      for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) H[r][c] += (object_points[i][r] - mo[r]) * (reprojected_points[i][c] - mr[c]);
    double A[3][3], V[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { A[r][c] = 0; for (int k = 0; k < 3; ++k) A[r][c] += H[k][r] * H[k][c]; }
    for (int sweep = 0; sweep < 60; ++sweep) {
      const double off = std::fabs(A[0][1]) + std::fabs(A[0][2]) + std::fabs(A[1][2]);
      if (off <= 1e-300 || off <= 1e-17 * (std::fabs(A[0][0]) + std::fabs(A[1][1]) + std::fabs(A[2][2]))) break;
      for (int p = 0; p < 2; ++p)
        for (int q = p + 1; q < 3; ++q) {
          if (A[p][q] == 0) continue;
          const double theta = (A[q][q] - A[p][p]) / (2 * A[p][q]);
          const double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1));
          const double c = 1 / std::sqrt(t * t + 1), sn = t * c;
          for (int k = 0; k < 3; ++k) { const double akp = A[k][p], akq = A[k][q]; A[k][p] = c * akp - sn * akq; A[k][q] = sn * akp + c * akq; }
          for (int k = 0; k < 3; ++k) { const double apk = A[p][k], aqk = A[q][k]; A[p][k] = c * apk - sn * aqk; A[q][k] = sn * apk + c * aqk; }
          for (int k = 0; k < 3; ++k) { const double vkp = V[k][p], vkq = V[k][q]; V[k][p] = c * vkp - sn * vkq; V[k][q] = sn * vkp + c * vkq; }
        }
    }
    int ord[3] = {0, 1, 2};
    for (int a = 0; a < 2; ++a) for (int b = a + 1; b < 3; ++b) if (A[ord[b]][ord[b]] > A[ord[a]][ord[a]]) std::swap(ord[a], ord[b]);
    double v[3][3], u[3][3], sig[3];   // v[i], u[i]: i-th right / left singular vector
    for (int i = 0; i < 3; ++i) {
      for (int k = 0; k < 3; ++k) v[i][k] = V[k][ord[i]];
      sig[i] = std::sqrt(std::max(0.0, A[ord[i]][ord[i]]));
    }
    for (int i = 0; i < 3; ++i) {
      if (i < 2 || sig[2] > 1e-7 * sig[0]) {   // singular values from H^T H carry half the digits
        for (int r = 0; r < 3; ++r) { u[i][r] = 0; for (int k = 0; k < 3; ++k) u[i][r] += H[r][k] * v[i][k]; u[i][r] /= sig[i]; }
      } else {   // rank 2 (coplanar points): complete both bases right-handed, which makes R a proper rotation
        u[2][0] = u[0][1] * u[1][2] - u[0][2] * u[1][1]; u[2][1] = u[0][2] * u[1][0] - u[0][0] * u[1][2]; u[2][2] = u[0][0] * u[1][1] - u[0][1] * u[1][0];
        v[2][0] = v[0][1] * v[1][2] - v[0][2] * v[1][1]; v[2][1] = v[0][2] * v[1][0] - v[0][0] * v[1][2]; v[2][2] = v[0][0] * v[1][1] - v[0][1] * v[1][0];
      }
    }
    Matrix4d T; T.fill(0.0); T[15] = 1.0;
    double R[3][3];
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) { R[r][c] = 0; for (int i = 0; i < 3; ++i) R[r][c] += v[i][r] * u[i][c]; }
    for (int r = 0; r < 3; ++r) {
      for (int c = 0; c < 3; ++c) T[4 * c + r] = R[r][c];
      T[12 + r] = mr[r] - (R[r][0] * mo[0] + R[r][1] * mo[1] + R[r][2] * mo[2]);
    }
    return T;
  }

  // initialise (:1503-1786) on the image points of setImagePoints(): device vote -> correspondencesFromHistogram ->
  // checkCorrespondences for every candidate set (device: pfmpe_p3p_check; host: the accept logic of :1427-1488) ->
  // particle seeds (:1432-1438, :1754-1758).  Returns 1 on success; predicted_pose_ / correspondences_ then hold the
  // first accepted set's pose and pairs, PoseParticle the seeds.
  unsigned initialise(int objectNumber)
  {
    if (!ensure() || objectNumber < 0 || objectNumber >= built_objects_) { setFlag(objectNumber, 13.f); return 0; }
    sync_params();
    const List4DPoints& obj = object_points_Vec[objectNumber];
    const unsigned L = (unsigned)obj.size(), D = (unsigned)image_points_.size();
    if (D < L) { setFlag(objectNumber, 10.f); return 0; }                                   // :1507-1512
    std::vector<uint32_t> hist_corr;
    if (!correspondenceHistogram(objectNumber, image_points_, hist_corr)) { setFlag(objectNumber, 13.f); return 0; }
    bool all_zero = true;
    for (size_t i = 0; i < hist_corr.size(); ++i) all_zero &= hist_corr[i] == 0;
    if (all_zero) { setFlag(objectNumber, 12.f); return 0; }                               // :1780
    const std::vector<VectorXuPairs> correspondencesVector = correspondencesFromHistogram(hist_corr, D, L, true);
    if (correspondencesVector.empty()) setFlag(objectNumber, 11.f);                        // :1723
    // checkCorrespondences for every candidate set (:1733-1751): the per-combination bodies run on the device, a chunk
    // of candidate sets per call; the loop ends early once every particle is seeded (no later set can be accepted:
    // :1741 needs NumberOfP3PEstimation < N_Particle), which only changes which failure flag is left behind
    const int M = (int)correspondencesVector.size();
    const int stride = (int)(L * (L - 1) * (L - 2) / 6), chunk_sets = 16384;
    std::vector<double> pts(2 * (size_t)D);
    for (unsigned i = 0; i < D; ++i) { pts[2 * i] = image_points_[i][0]; pts[2 * i + 1] = image_points_[i][1]; }
    PoseParticle_.assign((size_t)N_Particle, identity());
    int foundSol = 0, NumberOfP3PEstimation = 1, firstMatch = 0;
    VectorXuPairs correspondences_optimisation;
    Matrix4d predicted_pose_optimisation = identity();
    std::vector<uint32_t> rows;
    std::vector<int32_t> counts;
    std::vector<pfmpe_p3p_check_result> chk;
    for (int s0 = 0; s0 < M && NumberOfP3PEstimation <= N_Particle; s0 += chunk_sets) {
      const int m = std::min(chunk_sets, M - s0);
      rows.assign((size_t)m * 2 * PFMPE_MAX_LED, 0u);
      counts.assign(m, 0);
      for (int s = 0; s < m; ++s) {
        const VectorXuPairs& c = correspondencesVector[s0 + s];
        counts[s] = (int32_t)c.size();
        for (size_t i = 0; i < c.size() && i < PFMPE_MAX_LED; ++i) { rows[((size_t)s * PFMPE_MAX_LED + i) * 2] = c[i].first; rows[((size_t)s * PFMPE_MAX_LED + i) * 2 + 1] = c[i].second; }
      }
      chk.resize((size_t)m * stride);
      if (!check(pfmpe_p3p_check(h_, objectNumber, pts.data(), (int32_t)D, rows.data(), counts.data(), m, stride, certainty_threshold_, chk.data()))) {
        setFlag(objectNumber, 13.f);
        return 0;
      }
      for (int s = 0; s < m && NumberOfP3PEstimation <= N_Particle; ++s) {
        correspondences_ = correspondencesVector[s0 + s];
        if (checkCorrespondences(&chk[(size_t)s * stride], NumberOfP3PEstimation, objectNumber, (int)L) == 1 && NumberOfP3PEstimation < N_Particle) {
          foundSol = 1;
          if (firstMatch == 0) { firstMatch++; correspondences_optimisation = correspondences_; predicted_pose_optimisation = predicted_pose_; }
        }
      }
    }
    const int NumOfReasonableEstimations = NumberOfP3PEstimation - 1;                      // :1753-1758
    while (NumberOfP3PEstimation < N_Particle && foundSol == 1) {
      PoseParticle_[(size_t)(N_Particle - NumberOfP3PEstimation)] = PoseParticle_[(size_t)(N_Particle - NumberOfP3PEstimation + NumOfReasonableEstimations)];
      NumberOfP3PEstimation++;
    }
    correspondences_ = correspondences_optimisation;
    predicted_pose_ = predicted_pose_optimisation;
    return (unsigned)foundSol;
  }

  // The initialisation branch of estimateBodyPose (:149-187) for one object: setImagePoints, initialise(), then
  // current_pose_ = predicted_pose_, optimiseAndUpdatePose, PoseParticle_Vec = newPoseEstimation_Vec = PoseParticle,
  // it_since_initialized_ = 1, Flag_Fail = 0.  The next estimateBodyPose() runs the particle filter from there.
  bool initialiseBodyPose(int objectNumber, const List2DPoints& detected_led_positions, double time_to_predict)
  {
    if (objectNumber < 0 || objectNumber >= (int)object_points_Vec.size()) return false;
    if (detected_led_positions.size() < object_points_Vec[objectNumber].size()) { setFlag(objectNumber, 3.f); return false; }   // :189
    setImagePoints(detected_led_positions);
    if (initialise(objectNumber) != 1) return false;
    const Matrix4d unoptimised = predicted_pose_;                                          // current_pose_ = predicted_pose_ (:160)
    const List4DPoints& obj = object_points_Vec[objectNumber];
    std::vector<double> xyz(3 * obj.size()), pts(2 * image_points_.size());
    for (size_t j = 0; j < obj.size(); ++j) for (int c = 0; c < 3; ++c) xyz[3 * j + c] = obj[j][c];
    for (size_t i = 0; i < image_points_.size(); ++i) { pts[2 * i] = image_points_[i][0]; pts[2 * i + 1] = image_points_[i][1]; }
    std::vector<uint32_t> corr(2 * correspondences_.size());
    for (size_t i = 0; i < correspondences_.size(); ++i) { corr[2 * i] = correspondences_[i].first; corr[2 * i + 1] = correspondences_[i].second; }
    Matrix4d pose = predicted_pose_;
    Matrix6d cov;
    int iters = 0;
    if (!check(pfmpe_optimise_pose(camera_matrix_K_.data(), xyz.data(), (int)obj.size(), pts.data(), (int)image_points_.size(), corr.data(),
                                   (int)correspondences_.size(), pose.data(), cov.data(), &iters))) { setFlag(objectNumber, 13.f); return false; }
    predicted_pose_ = pose;                                                                // optimisePose (:1805-2009)
    std::vector<double> flat((size_t)N_Particle * 16);
    for (int n = 0; n < N_Particle; ++n) for (int i = 0; i < 16; ++i) flat[(size_t)n * 16 + i] = PoseParticle_[(size_t)n][i];
    const double t_prev = current_time_;                                                   // updatePose (:2011-2021)
    if (time_to_predict - current_time_ > 0.001 || time_to_predict < current_time_) current_time_ = time_to_predict;
    if (!check(pfmpe_set_particles(h_, objectNumber, flat.data(), N_Particle)) ||
        !check(pfmpe_track_set_state(h_, objectNumber, unoptimised.data(), t_prev, pose.data(), current_time_, 1u))) {
      setFlag(objectNumber, 13.f);
      return false;
    }
    results_[objectNumber] = pfmpe_track_result();
    for (int i = 0; i < 16; ++i) results_[objectNumber].pose[i] = pose[i];
    for (int i = 0; i < 36; ++i) results_[objectNumber].cov[i] = cov[i];
    PubData.numIter = iters;
    setFlag(objectNumber, 0.f);                                                            // :178
    return true;
  }
  void setImagePoints(List2DPoints points) { image_points_ = points; }                     // :929
  std::vector<Matrix4d> getSeededParticles() { return PoseParticle_; }
  VectorXuPairs getCorrespondencesOfInitialisation() { return correspondences_; }

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
  static Matrix4d identity() { Matrix4d m; m.fill(0.0); m[0] = m[5] = m[10] = m[15] = 1.0; return m; }
  void setFlag(int objectNumber, float f) { if (objectNumber >= 0 && objectNumber < 2) PubData.Flag_Fail[objectNumber] = f; }
  // checkCorrespondences (:1312-1501) for correspondences_, given the device's per-combination records
  unsigned checkCorrespondences(const pfmpe_p3p_check_result* rec, int& NumberOfP3PEstimation, int objectNumber, int minNumCorr)
  {
    const int n = (int)correspondences_.size();
    if (n < minNumCorr) { setFlag(objectNumber, 6.f); return 0; }                          // :1317-1321
    const List4DPoints& obj = object_points_Vec[objectNumber];
    const unsigned N = (unsigned)(n * (n - 1) * (n - 2) / 6);
    unsigned num_valid_correspondences = 0;
    std::vector<std::array<double, 3> > mean(obj.size());
    for (size_t j = 0; j < obj.size(); ++j) mean[j].fill(0.0);
    for (unsigned i = 0; i < N; ++i) {
      const pfmpe_p3p_check_result& r = rec[i];
      if (r.rc != 0) { setFlag(objectNumber, 9.f); continue; }                             // :1456
      if (!r.valid || r.best < 0) continue;
      num_valid_correspondences++;
      if (N_Particle >= NumberOfP3PEstimation) {                                           // :1430-1438
        Matrix4d m;
        for (int e = 0; e < 16; ++e) m[e] = r.pose[e];
        PoseParticle_[(size_t)(N_Particle - NumberOfP3PEstimation)] = m;
        NumberOfP3PEstimation++;
      }
      for (size_t j = 0; j < obj.size(); ++j)                                              // :1447-1451
        for (int k = 0; k < 3; ++k) mean[j][k] += r.pose[k] * obj[j][0] + r.pose[4 + k] * obj[j][1] + r.pose[8 + k] * obj[j][2] + r.pose[12 + k];
    }
    if ((double)num_valid_correspondences / N >= valid_correspondence_threshold_) {        // :1469-1480
      std::vector<std::array<double, 3> > o(obj.size());
      for (size_t j = 0; j < obj.size(); ++j) for (int k = 0; k < 3; ++k) { o[j][k] = obj[j][k]; mean[j][k] /= num_valid_correspondences; }
      predicted_pose_ = computeTransformation(o, mean);
      return 1;
    }
    setFlag(objectNumber, num_valid_correspondences > 0 ? 7.f : 8.f);                      // :1483-1486
    return 0;
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
  double certainty_threshold_ = 1.0, valid_correspondence_threshold_ = 0.5;   // cfg defaults (PFMonocularPoseEstimator.cfg:20-21)
  double current_time_ = 0.0;
  Matrix4d predicted_pose_ = identity();
  VectorXuPairs correspondences_;
  std::vector<Matrix4d> PoseParticle_;
  bool have_seed_;
  uint32_t seed_, frame_ = 0;
  std::random_device rd_;
  std::string err_;
};

}  // namespace pfmpe
#endif /* PFMPE_POSE_ESTIMATOR_HPP_ */
