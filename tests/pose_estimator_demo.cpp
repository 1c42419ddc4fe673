// Drives include/pfmpe_pose_estimator.hpp (the C++ mirror of the reference's PoseEstimator interface) over a
// recorded sequence, the way MPENode drives the reference (pf_mpe/src/monocular_pose_estimator.cpp:128,222-232,295-305).
// Input (text, written by tests/test_pose_estimator_cpp.py): N_Particle n_frames n_leds seed | markers (n_leds x 3) |
// K (9) | pose0 (16) t0 | pose1 (16) t1 | per frame: time n_points, points (n x 2).
// Output: per frame "found flag_fail numIter" + the 16 doubles of getPredictedPose(0) + covariance trace.
#include <cstdio>
#include <cstdlib>

#include "pfmpe_pose_estimator.hpp"

int main(int argc, char** argv)
{
  if (argc < 2) { std::fprintf(stderr, "usage: %s sequence.txt\n", argv[0]); return 2; }
  std::FILE* f = std::fopen(argv[1], "r");
  if (!f) { std::perror(argv[1]); return 2; }
  int N, F, L;
  unsigned seed;
  if (std::fscanf(f, "%d %d %d %u", &N, &F, &L, &seed) != 4) return 2;
  pfmpe::PoseEstimator trackable_object_;                       // the member MPENode holds (monocular_pose_estimator.h)
  pfmpe::List4DPoints markers(L);
  for (int j = 0; j < L; ++j) {
    if (std::fscanf(f, "%lf %lf %lf", &markers[j][0], &markers[j][1], &markers[j][2]) != 3) return 2;
    markers[j][3] = 1.0;
  }
  trackable_object_.setMarkerPositions(markers, std::vector<pfmpe::List4DPoints>(1, markers));        // :128
  for (int i = 0; i < 9; ++i) if (std::fscanf(f, "%lf", &trackable_object_.camera_matrix_K_[i]) != 1) return 2;   // :222-231
  trackable_object_.N_Particle = N;                                                                   // :498
  trackable_object_.setBackProjectionPixelTolerance(5.0);
  trackable_object_.back_projection_pixel_tolerance_PF = 4.0;
  trackable_object_.setSeed(seed);
  pfmpe::Matrix4d p0, p1;
  double t0, t1;
  for (int i = 0; i < 16; ++i) if (std::fscanf(f, "%lf", &p0[i]) != 1) return 2;
  if (std::fscanf(f, "%lf", &t0) != 1) return 2;
  for (int i = 0; i < 16; ++i) if (std::fscanf(f, "%lf", &p1[i]) != 1) return 2;
  if (std::fscanf(f, "%lf", &t1) != 1) return 2;
  const bool init_ok = trackable_object_.setInitialPoses(0, p0, t0, p1, t1);
  if (!init_ok) std::printf("init failed: %s\n", trackable_object_.lastError().c_str());
  for (int k = 0; k < F; ++k) {
    double t;
    int n;
    if (std::fscanf(f, "%lf %d", &t, &n) != 2) return 2;
    pfmpe::List2DPoints pts(n);
    for (int i = 0; i < n; ++i) if (std::fscanf(f, "%lf %lf", &pts[i][0], &pts[i][1]) != 2) return 2;
    const std::vector<bool> found = trackable_object_.estimateBodyPose(std::vector<pfmpe::List2DPoints>(1, pts), t);   // :295
    const pfmpe::Matrix4d pose = trackable_object_.getPredictedPose(0);                                                // :305
    const pfmpe::Matrix6d cov = trackable_object_.getPoseCovariance(0);                                                // :304
    const pfmpe::PoseEstimator::PubData_struct pd = trackable_object_.getPublisherData();                              // :418
    std::printf("%d %.1f %d", found[0] ? 1 : 0, pd.Flag_Fail[0], (int)pd.numIter);
    for (int i = 0; i < 16; ++i) std::printf(" %.17g", pose[i]);
    double tr = 0;
    for (int i = 0; i < 6; ++i) tr += cov[7 * i];
    std::printf(" %.17g\n", tr);
  }
  if (!trackable_object_.ok()) std::printf("no device: %s\n", trackable_object_.lastError().c_str());
  else {
    std::printf("particles %zu resampled %zu weights %zu correspondences %zu\n", trackable_object_.getPoseParticles(0).size(),
                trackable_object_.getResampledParticles(0).size(), trackable_object_.getParticleWeights(0).size(),
                trackable_object_.getCorrespondences().size());
  }
  std::fclose(f);
  return 0;
}
