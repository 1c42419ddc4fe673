/*
 * initialise_host.hpp -- ORACLE (test infrastructure, not product code).
 *
 * CPU restatement of the sequential HOST tail of PoseEstimator::initialise
 * (/root/reference/pf_mpe_lib/src/pose_estimator.cpp): correspondencesFromHistogram (:1134-1288), checkAmbiguity
 * (:2447-2458), the accept logic of checkCorrespondences (:1427-1488), computeTransformation (:2139-2161) and the
 * cyclic particle fill of initialise (:1726-1766).  SURVEY section 2 #8 marks this logic OUT OF SCOPE for the product:
 * the reference keeps these functions unchanged and only their two data-parallel loops (the P3P vote and the
 * per-combination bodies of checkCorrespondences) are replaced by device calls (pfmpe_p3p_histogram, pfmpe_p3p_check).
 * This header exists so that tests/ can string those two device calls together the way the reference does and check
 * that the seeds they produce initialise a filter that tracks on.  Only tests/ may include it; nothing under
 * pf_monocular_pose_estimator_b200/ or include/ does.  Parity unpinned: the reference has no vectors for it
 * (tests compare against NumPy restatements of the same lines and against np.linalg.svd).
 */
#ifndef ORACLE_INITIALISE_HOST_HPP_
#define ORACLE_INITIALISE_HOST_HPP_

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <utility>
#include <vector>

namespace orc_init {

typedef std::array<double, 16> Matrix4d;                                  // column-major (Eigen::Matrix4d layout)
typedef std::vector<std::pair<unsigned, unsigned> > VectorXuPairs;        // rows (LED, detection), 1-based
typedef std::array<double, 3> Point3;

inline Matrix4d identity() { Matrix4d m; m.fill(0.0); m[0] = m[5] = m[10] = m[15] = 1.0; return m; }

// checkAmbiguity (:2447-2458): two LEDs claim the same detection (the reference also reads one element past the end of
// the vector, which is not reproduced)
inline bool checkAmbiguity(const std::vector<int>& correspondingDetections)
{
  for (size_t i = 0; i + 1 < correspondingDetections.size(); ++i)
    for (size_t j = correspondingDetections.size() - 1; j > i; --j)
      if (correspondingDetections[i] == correspondingDetections[j]) return true;
  return false;
}

// correspondencesFromHistogram (:1134-1288): candidate correspondence sets, most probable first.
// histogram: numRows x numCols (detections x LEDs), row-major.
inline std::vector<VectorXuPairs> correspondencesFromHistogram(const std::vector<uint32_t>& histogram, unsigned numRows,
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

// computeTransformation (:2139-2161): rigid transform between two 3 x n point sets (columns), R = V U^T of the SVD of
// the cross-covariance (here: Jacobi on H^T H; a vanishing third singular value is completed to a proper rotation)
inline Matrix4d computeTransformation(const std::vector<Point3>& object_points, const std::vector<Point3>& reprojected_points)
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

// one per-combination record of checkCorrespondences' loop body (:1340-1445), as the device call returns it
struct CheckRecord {
  int rc;            // computePoses' return value
  int valid;         // valid_correspondence_found (:1412)
  int best;          // smallest_error_idx (:1419), -1 if none
  double pose[16];   // temp_solution.inverse() (:1437), column-major
};

// State of initialise()'s loop over the candidate sets (:1726-1766)
struct InitState {
  int N_Particle;
  std::vector<Matrix4d> PoseParticle;
  int NumberOfP3PEstimation;
  float flag;                        // last PubData.Flag_Fail value the accept logic wrote (0 = none)
  Matrix4d predicted_pose;
  explicit InitState(int n) : N_Particle(n), PoseParticle((size_t)n, identity()), NumberOfP3PEstimation(1), flag(0.f), predicted_pose(identity()) {}
};

// The accept logic of checkCorrespondences (:1427-1488) for one candidate set with n pairs, given the records of its
// C(n, 3) combinations.  Returns 1 if the set is accepted (st.predicted_pose then holds computeTransformation's result).
inline unsigned acceptCorrespondences(const CheckRecord* rec, int n, int minNumCorr, const std::vector<Point3>& obj,
                                      double valid_correspondence_threshold, InitState& st)
{
  if (n < minNumCorr) { st.flag = 6.f; return 0; }                                    // :1317-1321
  const unsigned N = (unsigned)(n * (n - 1) * (n - 2) / 6);
  unsigned num_valid_correspondences = 0;
  std::vector<Point3> mean(obj.size());
  for (size_t j = 0; j < obj.size(); ++j) mean[j].fill(0.0);
  for (unsigned i = 0; i < N; ++i) {
    const CheckRecord& r = rec[i];
    if (r.rc != 0) { st.flag = 9.f; continue; }                                       // :1456
    if (!r.valid || r.best < 0) continue;
    num_valid_correspondences++;
    if (st.N_Particle >= st.NumberOfP3PEstimation) {                                  // :1430-1438
      Matrix4d m;
      for (int e = 0; e < 16; ++e) m[e] = r.pose[e];
      st.PoseParticle[(size_t)(st.N_Particle - st.NumberOfP3PEstimation)] = m;
      st.NumberOfP3PEstimation++;
    }
    for (size_t j = 0; j < obj.size(); ++j)                                           // :1447-1451
      for (int k = 0; k < 3; ++k) mean[j][k] += r.pose[k] * obj[j][0] + r.pose[4 + k] * obj[j][1] + r.pose[8 + k] * obj[j][2] + r.pose[12 + k];
  }
  if ((double)num_valid_correspondences / N >= valid_correspondence_threshold) {      // :1469-1480
    for (size_t j = 0; j < obj.size(); ++j) for (int k = 0; k < 3; ++k) mean[j][k] /= num_valid_correspondences;
    st.predicted_pose = computeTransformation(obj, mean);
    return 1;
  }
  st.flag = num_valid_correspondences > 0 ? 7.f : 8.f;                                // :1483-1486
  return 0;
}

// cyclic fill of the particles no P3P solution seeded (:1753-1758)
inline void fillRemainingParticles(InitState& st, int foundSol)
{
  const int NumOfReasonableEstimations = st.NumberOfP3PEstimation - 1;
  while (st.NumberOfP3PEstimation < st.N_Particle && foundSol == 1) {
    st.PoseParticle[(size_t)(st.N_Particle - st.NumberOfP3PEstimation)] =
        st.PoseParticle[(size_t)(st.N_Particle - st.NumberOfP3PEstimation + NumOfReasonableEstimations)];
    st.NumberOfP3PEstimation++;
  }
}

}  // namespace orc_init
#endif /* ORACLE_INITIALISE_HOST_HPP_ */
