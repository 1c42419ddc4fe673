// Drives the initialisation side: the reference's host tail of initialise() (restated in oracle/initialise_host.hpp:
// test infrastructure) strung together with the two device calls of the C++ mirror (include/pfmpe_pose_estimator.hpp:
// correspondenceHistogram, checkCorrespondenceSets) the way INTEGRATION.md section 5 wires them into the reference.
// Modes (first argument):
//   hist  <file>: numRows numCols, then the histogram row-major -> the candidate sets of correspondencesFromHistogram
//   xform <file>: n, then n object points (3) and n reprojected points (3) -> computeTransformation (16, column-major)
//   init  <file>: N_Particle n_leds seed n_objects | markers | K (9) | t_init n, points | n_frames | per frame: time n, points
//                 -> initialisation of every object (same detections) + the tracked frames (found flag_fail, pose of
//                 object 0, then of the last object); needs a CUDA device
#include <chrono>
#include <cstdio>
#include <cstring>

#include "../oracle/initialise_host.hpp"
#include "pfmpe_pose_estimator.hpp"

static int fail(const char* what) { std::fprintf(stderr, "bad input: %s\n", what); return 2; }

// PoseEstimator::initialise (:1503-1786) + the branch of estimateBodyPose that calls it (:149-187) for one object:
// device vote -> correspondencesFromHistogram -> device checkCorrespondences bodies -> accept logic -> seeds ->
// optimisePose -> filter state.
static bool initialise_object(pfmpe::PoseEstimator& est, int objectNumber, const pfmpe::List4DPoints& markers, int N_Particle,
                              const pfmpe::List2DPoints& pts, double t, pfmpe::VectorXuPairs& corr_out, float& flag)
{
  const unsigned L = (unsigned)markers.size(), D = (unsigned)pts.size();
  flag = 0.f;
  if (D < L) { flag = 10.f; return false; }                                              // :1507-1512
  std::vector<uint32_t> hist;
  if (!est.correspondenceHistogram(objectNumber, pts, hist)) { flag = pfmpe::kFlagFailNoDevice; return false; }
  bool all_zero = true;
  for (size_t i = 0; i < hist.size(); ++i) all_zero = all_zero && hist[i] == 0;
  if (all_zero) { flag = 12.f; return false; }                                           // :1780
  const std::vector<orc_init::VectorXuPairs> sets = orc_init::correspondencesFromHistogram(hist, D, L, true);
  if (sets.empty()) { flag = 11.f; return false; }                                       // :1723
  std::vector<pfmpe_p3p_check_result> rec;
  int stride = 0;
  if (!est.checkCorrespondenceSets(objectNumber, pts, sets, rec, stride)) { flag = pfmpe::kFlagFailNoDevice; return false; }
  std::vector<orc_init::Point3> obj(L);
  for (unsigned j = 0; j < L; ++j) for (int k = 0; k < 3; ++k) obj[j][k] = markers[j][k];
  orc_init::InitState st(N_Particle);
  int foundSol = 0, firstMatch = 0;
  orc_init::Matrix4d first_pose = orc_init::identity();
  std::vector<orc_init::CheckRecord> one((size_t)stride);
  for (size_t s = 0; s < sets.size() && st.NumberOfP3PEstimation <= N_Particle; ++s) {   // :1733-1751
    for (int c = 0; c < stride; ++c) {
      const pfmpe_p3p_check_result& r = rec[s * (size_t)stride + c];
      one[c].rc = r.rc; one[c].valid = r.valid; one[c].best = r.best;
      std::memcpy(one[c].pose, r.pose, sizeof one[c].pose);
    }
    if (orc_init::acceptCorrespondences(one.data(), (int)sets[s].size(), (int)L, obj, 0.5, st) == 1 && st.NumberOfP3PEstimation < N_Particle) {
      foundSol = 1;
      if (firstMatch == 0) { firstMatch = 1; corr_out = sets[s]; first_pose = st.predicted_pose; }
    }
  }
  orc_init::fillRemainingParticles(st, foundSol);                                        // :1753-1758
  if (!foundSol) { flag = st.flag; return false; }
  pfmpe::Matrix4d unoptimised = first_pose, pose = first_pose;                           // current_pose_ = predicted_pose_ (:160)
  pfmpe::Matrix6d cov;
  int iters = 0;
  if (!est.optimisePose(objectNumber, pts, corr_out, pose, cov, iters)) { flag = pfmpe::kFlagFailNoDevice; return false; }
  std::vector<pfmpe::Matrix4d> seeds(st.PoseParticle.begin(), st.PoseParticle.end());
  if (!est.seedInitialisedObject(objectNumber, seeds, unoptimised, pose, cov, iters, t)) { flag = pfmpe::kFlagFailNoDevice; return false; }
  return true;
}

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
    const std::vector<orc_init::VectorXuPairs> v = orc_init::correspondencesFromHistogram(h, R, C, true);
    for (size_t s = 0; s < v.size(); ++s) {
      for (size_t i = 0; i < v[s].size(); ++i) std::printf("%u:%u ", v[s][i].first, v[s][i].second);
      std::printf("\n");
    }
    return 0;
  }
  if (!std::strcmp(argv[1], "xform")) {
    int n;
    if (std::fscanf(f, "%d", &n) != 1) return fail("n");
    std::vector<orc_init::Point3> a(n), b(n);
    for (int i = 0; i < n; ++i) if (std::fscanf(f, "%lf %lf %lf", &a[i][0], &a[i][1], &a[i][2]) != 3) return fail("a");
    for (int i = 0; i < n; ++i) if (std::fscanf(f, "%lf %lf %lf", &b[i][0], &b[i][1], &b[i][2]) != 3) return fail("b");
    const orc_init::Matrix4d T = orc_init::computeTransformation(a, b);
    for (int i = 0; i < 16; ++i) std::printf("%.17g ", T[i]);
    std::printf("\n");
    return 0;
  }
  if (!std::strcmp(argv[1], "init")) {
    int N, L, n, F, T;
    unsigned seed;
    double t;
    if (std::fscanf(f, "%d %d %u %d", &N, &L, &seed, &T) != 4) return fail("header");
    pfmpe::PoseEstimator est;
    pfmpe::List4DPoints markers(L);
    for (int j = 0; j < L; ++j) { if (std::fscanf(f, "%lf %lf %lf", &markers[j][0], &markers[j][1], &markers[j][2]) != 3) return fail("markers"); markers[j][3] = 1; }
    est.setMarkerPositions(markers, std::vector<pfmpe::List4DPoints>(T, markers));
    for (int i = 0; i < 9; ++i) if (std::fscanf(f, "%lf", &est.camera_matrix_K_[i]) != 1) return fail("K");
    est.N_Particle = N;
    est.setSeed(seed);
    if (std::fscanf(f, "%lf %d", &t, &n) != 2) return fail("init frame");
    pfmpe::List2DPoints pts(n);
    for (int i = 0; i < n; ++i) if (std::fscanf(f, "%lf %lf", &pts[i][0], &pts[i][1]) != 2) return fail("points");
    pfmpe::VectorXuPairs c;
    float flag = 0.f;
    bool ok = true;
    for (int o = 0; o < T; ++o) ok = initialise_object(est, o, markers, N, pts, t, c, flag) && ok;   // every object at the same timestamp
    if (argc > 3) {   // timing: repeat (the first call also creates the handle and the CUDA context)
      for (int rep = 0; rep < 3; ++rep) {
        const auto a = std::chrono::steady_clock::now();
        ok = initialise_object(est, 0, markers, N, pts, t, c, flag);
        std::fprintf(stderr, "initialisation: %.3f ms (%d detections)\n",
                     1e3 * std::chrono::duration<double>(std::chrono::steady_clock::now() - a).count(), n);
      }
    }
    std::printf("init %d flag %.1f", ok ? 1 : 0, ok ? est.getPublisherData().Flag_Fail[0] : flag);
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
      const std::vector<bool> found = est.estimateBodyPose(std::vector<pfmpe::List2DPoints>(T, pts), t);
      const pfmpe::Matrix4d p = est.getPredictedPose(0), q = est.getPredictedPose(T - 1);
      std::printf("%d %.1f", (found[0] && found[T - 1]) ? 1 : 0, est.getPublisherData().Flag_Fail[0]);
      for (int i = 0; i < 16; ++i) std::printf(" %.17g", p[i]);
      for (int i = 0; i < 16; ++i) std::printf(" %.17g", q[i]);
      std::printf("\n");
    }
    return 0;
  }
  return fail("mode");
}
