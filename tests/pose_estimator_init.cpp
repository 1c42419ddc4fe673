// Drives the initialisation side of include/pfmpe_pose_estimator.hpp.  Modes (first argument):
//   hist  <file>: numRows numCols, then the histogram row-major -> the candidate sets of correspondencesFromHistogram
//   xform <file>: n, then n object points (3) and n reprojected points (3) -> computeTransformation (16, column-major)
//   init  <file>: N_Particle n_leds seed | markers | K (9) | t_init n, points | n_frames | per frame: time n, points
//                 -> initialiseBodyPose + the tracked frames (found flag_fail, pose) ; needs a CUDA device
#include <chrono>
#include <cstdio>
#include <cstring>

#include "pfmpe_pose_estimator.hpp"

static int fail(const char* what) { std::fprintf(stderr, "bad input: %s\n", what); return 2; }

int main(int argc, char** argv)
{
  if (argc < 3) return fail("usage: mode file");
  std::FILE* f = std::fopen(argv[2], "r");
  if (!f) return fail("cannot open");
  if (!std::strcmp(argv[1], "hist")) {
    unsigned R, C;
    if (std::fscanf(f, "%u %u", &R, &C) != 2) return fail("shape");
    std::vector<uint32_t> h((size_t)R * C);
    for (size_t i = 0; i < h.size(); ++i) if (std::fscanf(f, "%u", &h[i]) != 1) return fail("hist");
    const std::vector<pfmpe::VectorXuPairs> v = pfmpe::PoseEstimator::correspondencesFromHistogram(h, R, C, true);
    for (size_t s = 0; s < v.size(); ++s) {
      for (size_t i = 0; i < v[s].size(); ++i) std::printf("%u:%u ", v[s][i].first, v[s][i].second);
      std::printf("\n");
    }
    return 0;
  }
  if (!std::strcmp(argv[1], "xform")) {
    int n;
    if (std::fscanf(f, "%d", &n) != 1) return fail("n");
    std::vector<std::array<double, 3> > a(n), b(n);
    for (int i = 0; i < n; ++i) if (std::fscanf(f, "%lf %lf %lf", &a[i][0], &a[i][1], &a[i][2]) != 3) return fail("a");
    for (int i = 0; i < n; ++i) if (std::fscanf(f, "%lf %lf %lf", &b[i][0], &b[i][1], &b[i][2]) != 3) return fail("b");
    const pfmpe::Matrix4d T = pfmpe::PoseEstimator::computeTransformation(a, b);
    for (int i = 0; i < 16; ++i) std::printf("%.17g ", T[i]);
    std::printf("\n");
    return 0;
  }
  if (!std::strcmp(argv[1], "init")) {
    int N, L, n, F;
    unsigned seed;
    double t;
    if (std::fscanf(f, "%d %d %u", &N, &L, &seed) != 3) return fail("header");
    pfmpe::PoseEstimator est;
    pfmpe::List4DPoints markers(L);
    for (int j = 0; j < L; ++j) { if (std::fscanf(f, "%lf %lf %lf", &markers[j][0], &markers[j][1], &markers[j][2]) != 3) return fail("markers"); markers[j][3] = 1; }
    est.setMarkerPositions(markers, std::vector<pfmpe::List4DPoints>(1, markers));
    for (int i = 0; i < 9; ++i) if (std::fscanf(f, "%lf", &est.camera_matrix_K_[i]) != 1) return fail("K");
    est.N_Particle = N;
    est.setSeed(seed);
    if (std::fscanf(f, "%lf %d", &t, &n) != 2) return fail("init frame");
    pfmpe::List2DPoints pts(n);
    for (int i = 0; i < n; ++i) if (std::fscanf(f, "%lf %lf", &pts[i][0], &pts[i][1]) != 2) return fail("points");
    bool ok = est.initialiseBodyPose(0, pts, t);
    if (argc > 3) {   // timing: repeat (the first call also creates the handle and the CUDA context)
      for (int rep = 0; rep < 3; ++rep) {
        const auto a = std::chrono::steady_clock::now();
        ok = est.initialiseBodyPose(0, pts, t);
        std::fprintf(stderr, "initialiseBodyPose: %.3f ms (%d detections)\n",
                     1e3 * std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count(), n);
      }
    }
    std::printf("init %d flag %.1f", ok ? 1 : 0, est.getPublisherData().Flag_Fail[0]);
    const pfmpe::VectorXuPairs c = est.getCorrespondencesOfInitialisation();
    for (size_t i = 0; i < c.size(); ++i) std::printf(" %u:%u", c[i].first, c[i].second);
    std::printf("\n");
    const pfmpe::Matrix4d p0 = est.getPredictedPose(0);
    std::printf("pose");
    for (int i = 0; i < 16; ++i) std::printf(" %.17g", p0[i]);
    std::printf("\n");
    if (!ok) { std::printf("error: %s\n", est.lastError().c_str()); return 0; }
    if (std::fscanf(f, "%d", &F) != 1) return fail("n_frames");
    for (int k = 0; k < F; ++k) {
      if (std::fscanf(f, "%lf %d", &t, &n) != 2) return fail("frame");
      pts.resize(n);
      for (int i = 0; i < n; ++i) if (std::fscanf(f, "%lf %lf", &pts[i][0], &pts[i][1]) != 2) return fail("points");
      const std::vector<bool> found = est.estimateBodyPose(std::vector<pfmpe::List2DPoints>(1, pts), t);
      const pfmpe::Matrix4d p = est.getPredictedPose(0);
      std::printf("%d %.1f", found[0] ? 1 : 0, est.getPublisherData().Flag_Fail[0]);
      for (int i = 0; i < 16; ++i) std::printf(" %.17g", p[i]);
      std::printf("\n");
    }
    return 0;
  }
  return fail("mode");
}
