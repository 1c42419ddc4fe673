/*
 * p3p_oracle.cpp -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * C++11 restatement, in double precision, of the brute-force (re-)initialisation vote of pf_mpe_lib
 * (SURVEY 8 row f4):
 *   - PoseEstimator::calculateImageVectors      /root/reference/pf_mpe_lib/src/pose_estimator.cpp:1072-1085
 *   - P3P::computePoses / P3P::solveQuartic     /root/reference/pf_mpe_lib/src/p3p.cpp:65-236, :238-292
 *   - PoseEstimator::calculateMinDistancesAndPairs                        pose_estimator.cpp:2093-2137
 *   - the histogram loop of PoseEstimator::initialise                     pose_estimator.cpp:1524-1717
 * std::complex is used where the reference uses it, so the quartic follows libstdc++'s rules for
 * pow(complex, double) / sqrt(complex) exactly as a build of the reference with this toolchain would.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may call this.
 *
 * PARITY PINNING STATUS: "parity unpinned" -- the reference has no tests / golden vectors for this path and
 * cannot be compiled here (Eigen, OpenCV, ROS absent).  Eigen's evaluation order of the small fixed-size
 * products is restated as plain left-to-right sums.
 *
 * Layouts: 3x3 matrices are column-major double[9] (Eigen::Matrix3d); a P3P solution is a column-major 3x4
 * double[12] = [R | C]; image points double[2]; object points homogeneous double[4]; the histogram is
 * row-major [D][L] (hist_corr(image point, object point)).
 */
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstring>
#include <limits>
#include <vector>

extern "C" {
void orc_mat4_inv(const double *A, double *Ainv);
void orc_project2d(const double K[9], const double T[16], const double X[4], double uv[2]);
}

namespace {

struct V3 { double x, y, z; };
inline V3 sub(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 add(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 cross(V3 a, V3 b) { return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline double norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 divs(V3 a, double s) { return V3{a.x / s, a.y / s, a.z / s}; }
inline V3 col(const double *M, int c) { return V3{M[3 * c], M[3 * c + 1], M[3 * c + 2]}; }
/* rows r0,r1,r2 times v */
inline V3 rows_mul(V3 r0, V3 r1, V3 r2, V3 v) { return V3{dot(r0, v), dot(r1, v), dot(r2, v)}; }
/* transpose(rows) times v = r0*v.x + r1*v.y + r2*v.z */
inline V3 rowsT_mul(V3 r0, V3 r1, V3 r2, V3 v)
{
  return V3{r0.x * v.x + r1.x * v.y + r2.x * v.z, r0.y * v.x + r1.y * v.y + r2.y * v.z,
            r0.z * v.x + r1.z * v.y + r2.z * v.z};
}

/* P3P::solveQuartic (p3p.cpp:238-292) */
void solve_quartic(const double f[5], double roots[4])
{
  typedef std::complex<double> cx;
  const double A = f[0], B = f[1], C = f[2], D = f[3], E = f[4];
  const double A_pw2 = A * A, B_pw2 = B * B, A_pw3 = A_pw2 * A, B_pw3 = B_pw2 * B, A_pw4 = A_pw3 * A,
               B_pw4 = B_pw3 * B;
  const double alpha = -3 * B_pw2 / (8 * A_pw2) + C / A;
  const double beta = B_pw3 / (8 * A_pw3) - B * C / (2 * A_pw2) + D / A;
  const double gamma = -3 * B_pw4 / (256 * A_pw4) + B_pw2 * C / (16 * A_pw3) - B * D / (4 * A_pw2) + E / A;
  const double alpha_pw2 = alpha * alpha, alpha_pw3 = alpha_pw2 * alpha;

  cx P(-alpha_pw2 / 12 - gamma, 0);
  cx Q(-alpha_pw3 / 108 + alpha * gamma / 3 - std::pow(beta, 2) / 8, 0);
  cx R = -Q / 2.0 + std::sqrt(std::pow(Q, 2.0) / 4.0 + std::pow(P, 3.0) / 27.0);
  cx U = std::pow(R, (1.0 / 3.0));
  cx y;
  if (U.real() == 0)
    y = -5.0 * alpha / 6.0 - std::pow(Q, (1.0 / 3.0));
  else
    y = -5.0 * alpha / 6.0 - P / (3.0 * U) + U;
  cx w = std::sqrt(alpha + 2.0 * y);
  cx temp;
  temp = -B / (4.0 * A) + 0.5 * (w + std::sqrt(-(3.0 * alpha + 2.0 * y + 2.0 * beta / w)));
  roots[0] = temp.real();
  temp = -B / (4.0 * A) + 0.5 * (w - std::sqrt(-(3.0 * alpha + 2.0 * y + 2.0 * beta / w)));
  roots[1] = temp.real();
  temp = -B / (4.0 * A) + 0.5 * (-w + std::sqrt(-(3.0 * alpha + 2.0 * y - 2.0 * beta / w)));
  roots[2] = temp.real();
  temp = -B / (4.0 * A) + 0.5 * (-w - std::sqrt(-(3.0 * alpha + 2.0 * y - 2.0 * beta / w)));
  roots[3] = temp.real();
}

/* P3P::computePoses (p3p.cpp:65-236) */
int compute_poses(const double *fv, const double *wp, double *sol)
{
  V3 P1 = col(wp, 0), P2 = col(wp, 1), P3 = col(wp, 2);
  if (norm(cross(sub(P2, P1), sub(P3, P1))) == 0) return -1; /* :78-84 colinear world points */

  V3 f1 = col(fv, 0), f2 = col(fv, 1), f3 = col(fv, 2);
  V3 e1 = f1, e3 = cross(f1, f2);
  e3 = divs(e3, norm(e3));
  V3 e2 = cross(e3, e1);
  f3 = rows_mul(e1, e2, e3, f3); /* T * f3 */
  if (f3.z > 0) { /* :104-125 */
    f1 = col(fv, 1); f2 = col(fv, 0); f3 = col(fv, 2);
    e1 = f1; e3 = cross(f1, f2);
    e3 = divs(e3, norm(e3));
    e2 = cross(e3, e1);
    f3 = rows_mul(e1, e2, e3, f3);
    P1 = col(wp, 1); P2 = col(wp, 0); P3 = col(wp, 2);
  }
  V3 n1 = sub(P2, P1);
  n1 = divs(n1, norm(n1));
  V3 n3 = cross(n1, sub(P3, P1));
  n3 = divs(n3, norm(n3));
  V3 n2 = cross(n3, n1);
  P3 = rows_mul(n1, n2, n3, sub(P3, P1)); /* N * (P3 - P1) */

  const double d_12 = norm(sub(P2, P1));
  const double f_1 = f3.x / f3.z, f_2 = f3.y / f3.z, p_1 = P3.x, p_2 = P3.y;
  const double cos_beta = dot(f1, f2);
  double b = 1 / (1 - std::pow(cos_beta, 2)) - 1;
  b = (cos_beta < 0) ? -std::sqrt(b) : std::sqrt(b);

  const double f_1_pw2 = std::pow(f_1, 2), f_2_pw2 = std::pow(f_2, 2), p_1_pw2 = std::pow(p_1, 2);
  const double p_1_pw3 = p_1_pw2 * p_1, p_1_pw4 = p_1_pw3 * p_1, p_2_pw2 = std::pow(p_2, 2);
  const double p_2_pw3 = p_2_pw2 * p_2, p_2_pw4 = p_2_pw3 * p_2, d_12_pw2 = std::pow(d_12, 2),
               b_pw2 = std::pow(b, 2);
  double fac[5];
  fac[0] = -f_2_pw2 * p_2_pw4 - p_2_pw4 * f_1_pw2 - p_2_pw4;
  fac[1] = 2 * p_2_pw3 * d_12 * b + 2 * f_2_pw2 * p_2_pw3 * d_12 * b - 2 * f_2 * p_2_pw3 * f_1 * d_12;
  fac[2] = -f_2_pw2 * p_2_pw2 * p_1_pw2 - f_2_pw2 * p_2_pw2 * d_12_pw2 * b_pw2 - f_2_pw2 * p_2_pw2 * d_12_pw2 +
           f_2_pw2 * p_2_pw4 + p_2_pw4 * f_1_pw2 + 2 * p_1 * p_2_pw2 * d_12 +
           2 * f_1 * f_2 * p_1 * p_2_pw2 * d_12 * b - p_2_pw2 * p_1_pw2 * f_1_pw2 +
           2 * p_1 * p_2_pw2 * f_2_pw2 * d_12 - p_2_pw2 * d_12_pw2 * b_pw2 - 2 * p_1_pw2 * p_2_pw2;
  fac[3] = 2 * p_1_pw2 * p_2 * d_12 * b + 2 * f_2 * p_2_pw3 * f_1 * d_12 - 2 * f_2_pw2 * p_2_pw3 * d_12 * b -
           2 * p_1 * p_2 * d_12_pw2 * b;
  fac[4] = -2 * f_2 * p_2_pw2 * f_1 * p_1 * d_12 * b + f_2_pw2 * p_2_pw2 * d_12_pw2 + 2 * p_1_pw3 * d_12 -
           p_1_pw2 * d_12_pw2 + f_2_pw2 * p_2_pw2 * p_1_pw2 - p_1_pw4 - 2 * f_2_pw2 * p_2_pw2 * p_1 * d_12 +
           p_2_pw2 * f_1_pw2 * p_1_pw2 + f_2_pw2 * p_2_pw2 * d_12_pw2 * b_pw2;
  double roots[4];
  solve_quartic(fac, roots);

  for (int i = 0; i < 4; ++i) { /* :181-233 back-substitution */
    const double cot_alpha =
        (-f_1 * p_1 / f_2 - roots[i] * p_2 + d_12 * b) / (-f_1 * roots[i] * p_2 / f_2 + p_1 - d_12);
    const double cos_theta = roots[i];
    const double sin_theta = std::sqrt(1 - std::pow(roots[i], 2));
    const double sin_alpha = std::sqrt(1 / (std::pow(cot_alpha, 2) + 1));
    double cos_alpha = std::sqrt(1 - std::pow(sin_alpha, 2));
    if (cot_alpha < 0) cos_alpha = -cos_alpha;

    V3 C;
    C.x = d_12 * cos_alpha * (sin_alpha * b + cos_alpha);
    C.y = cos_theta * d_12 * sin_alpha * (sin_alpha * b + cos_alpha);
    C.z = sin_theta * d_12 * sin_alpha * (sin_alpha * b + cos_alpha);
    C = add(P1, rowsT_mul(n1, n2, n3, C));

    /* R' (row-major as written at :205-214); result = N^T * R'^T * T evaluated left to right */
    const double Rp[3][3] = {{-cos_alpha, -sin_alpha * cos_theta, -sin_alpha * sin_theta},
                             {sin_alpha, -cos_alpha * cos_theta, -cos_alpha * sin_theta},
                             {0, -sin_theta, cos_theta}};
    const double Nm[3][3] = {{n1.x, n1.y, n1.z}, {n2.x, n2.y, n2.z}, {n3.x, n3.y, n3.z}};
    const double Tm[3][3] = {{e1.x, e1.y, e1.z}, {e2.x, e2.y, e2.z}, {e3.x, e3.y, e3.z}};
    double M1[3][3], Rm[3][3];
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) { /* (N^T R'^T)(r,c) = sum_k N(k,r) R'(c,k) */
        double s = 0;
        for (int k = 0; k < 3; ++k) s += Nm[k][r] * Rp[c][k];
        M1[r][c] = s;
      }
    for (int r = 0; r < 3; ++r)
      for (int c = 0; c < 3; ++c) {
        double s = 0;
        for (int k = 0; k < 3; ++k) s += M1[r][k] * Tm[k][c];
        Rm[r][c] = s;
      }
    double *S = sol + 12 * i;
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) S[3 * c + r] = Rm[r][c];
    S[9] = C.x; S[10] = C.y; S[11] = C.z;
  }
  return 0;
}

inline double sqdist2(const double *a, const double *b)
{
  const double dx = a[0] - b[0], dy = a[1] - b[1];
  return dx * dx + dy * dy;
}

} // namespace

extern "C" {

int orc_p3p_solve_quartic(const double factors[5], double roots[4])
{
  solve_quartic(factors, roots);
  return 0;
}

int orc_p3p_compute_poses(const double *feature_vectors, const double *world_points, double *solutions)
{
  return compute_poses(feature_vectors, world_points, solutions);
}

/* calculateImageVectors (pose_estimator.cpp:1072-1085): vec is D x 3 */
void orc_image_vectors(const double K[9], int D, const double *img, double *vec)
{
  for (int i = 0; i < D; ++i) {
    V3 v{(img[2 * i] - K[2]) / K[0], (img[2 * i + 1] - K[5]) / K[4], 1.0};
    v = divs(v, norm(v));
    vec[3 * i] = v.x; vec[3 * i + 1] = v.y; vec[3 * i + 2] = v.z;
  }
}

/* calculateMinDistancesAndPairs (pose_estimator.cpp:2093-2137): a: na x 2, b: nb x 2;
 * pair_b[i] = 1-based index of the first nearest b (0 if none compares smaller than +INF). */
void orc_min_distances_and_pairs(int na, const double *a, int nb, const double *b, double *min_dist,
                                 unsigned *pair_b)
{
  for (int i = 0; i < na; ++i) {
    double best = std::numeric_limits<double>::infinity();
    unsigned bj = 0;
    for (int j = 0; j < nb; ++j) {
      const double d = sqdist2(a + 2 * i, b + 2 * j);
      if (d < best) { best = d; bj = (unsigned)j + 1; }
    }
    min_dist[i] = std::sqrt(best);
    pair_b[i] = bj;
  }
}

/* The vote of PoseEstimator::initialise (pose_estimator.cpp:1524-1717).
 * obj: L x 4 homogeneous, img: D x 2 undistorted pixels, tol = back_projection_pixel_tolerance_.
 * hist: D x L (row-major, zeroed here).  stats (optional, 4 entries): image-point triples that pass the
 * pre-filters (:1556-1581), P3P::computePoses calls that returned 0, solutions back-projected
 * (finite and not skipped at :1659-1663), solutions that voted (:1692).  Returns 0. */
int orc_p3p_histogram_mode(const double K[9], int L, const double *obj, int D, const double *img, double tol,
                           int dup_mode, unsigned *hist, uint64_t *stats);

int orc_p3p_histogram(const double K[9], int L, const double *obj, int D, const double *img, double tol,
                      unsigned *hist, uint64_t *stats)
{
  return orc_p3p_histogram_mode(K, L, obj, D, img, tol, 0, hist, stats);
}

/* dup_mode selects the duplicate-solution rule of :1659-1663, whose outcome for a complex-conjugate root pair
 * (equal real parts up to rounding) depends on the last bit of the libm in use:
 *   0  the reference's rule: drop solution k when any entry equals solution k-1's exactly;
 *   1  never drop (upper bracket: every vote any rounding could produce);
 *   2  drop when the reference's rule fires OR all entries agree within 1e-6 (lower bracket).
 * Votes are non-negative, so cell by cell  mode 2 <= mode 0 <= mode 1, and the same holds for any
 * implementation whose arithmetic differs from this one by rounding only. */
int orc_p3p_histogram_mode(const double K[9], int L, const double *obj, int D, const double *img, double tol,
                           int dup_mode, unsigned *hist, uint64_t *stats)
{
  std::memset(hist, 0, sizeof(unsigned) * (size_t)D * (size_t)L);
  uint64_t st[4] = {0, 0, 0, 0};
  std::vector<double> vec(3 * (size_t)D);
  orc_image_vectors(K, D, img, vec.data());
  const double threshDist = 10000 * 100, threshDist2 = 10000 * 100; /* :1558-1559 */
  std::vector<double> un_im(2 * (size_t)D), mind(D), proj(2 * (size_t)L);
  std::vector<unsigned> un_im_idx(D), pr(D);

  for (int i0 = 0; i0 < D; ++i0)
    for (int i1 = i0 + 1; i1 < D; ++i1)
      for (int i2 = i1 + 1; i2 < D; ++i2) {
        const int ci[3] = {i0, i1, i2};
        double fv[9];
        for (int c = 0; c < 3; ++c)
          for (int r = 0; r < 3; ++r) fv[3 * c + r] = vec[3 * ci[c] + r];
        const double *d1 = img + 2 * i0, *d2 = img + 2 * i1, *d3 = img + 2 * i2;
        if (sqdist2(d1, d2) > threshDist) continue;
        else if (sqdist2(d1, d3) > threshDist) continue;
        else if (sqdist2(d2, d3) > threshDist) continue;
        const double dm[2] = {(d1[0] + d2[0] + d3[0]) / 3, (d1[1] + d2[1] + d3[1]) / 3};
        int cd = 0;
        for (int kk = 0; kk < D; ++kk)
          if (sqdist2(dm, img + 2 * kk) < threshDist2) cd++;
        if (cd < 5) continue; /* :1578 */
        st[0]++;
        /* unused image points close to the centroid (:1583-1607) */
        const int total_unused_im = D - 3;
        int n_un_im = 0;
        for (int kk = 0; kk < D; ++kk) {
          const bool used = (kk == i0 || kk == i1 || kk == i2);
          if (!used && sqdist2(dm, img + 2 * kk) < threshDist2) {
            un_im[2 * n_un_im] = img[2 * kk];
            un_im[2 * n_un_im + 1] = img[2 * kk + 1];
            un_im_idx[n_un_im] = (unsigned)kk;
            n_un_im++;
          }
          if (n_un_im == total_unused_im) break;
        }
        for (int j0 = 0; j0 < L; ++j0)
          for (int j1 = 0; j1 < L; ++j1)
            for (int j2 = 0; j2 < L; ++j2) {
              if (j0 == j1 || j0 == j2 || j1 == j2) continue;
              const int pj[3] = {j0, j1, j2};
              double wp[9];
              for (int c = 0; c < 3; ++c)
                for (int r = 0; r < 3; ++r) wp[3 * c + r] = obj[4 * pj[c] + r];
              double sol[48];
              if (compute_poses(fv, wp, sol) != 0) continue;
              st[1]++;
              /* unused object points, ascending (:1637-1656) */
              int un_obj_idx[16], n_un_obj = 0;
              for (int ll = 0; ll < L; ++ll)
                if (ll != j0 && ll != j1 && ll != j2) un_obj_idx[n_un_obj++] = ll;
              for (int k = 0; k < 4; ++k) {
                if (k > 0) { /* :1659-1663: (sol(k) - sol(k-1)).all() == 0  <=>  some entry of the difference is 0 */
                  bool all_nonzero = true, all_close = true;
                  for (int e = 0; e < 12; ++e) {
                    const double df = sol[12 * k + e] - sol[12 * (k - 1) + e];
                    if (df == 0) all_nonzero = false;
                    if (!(std::fabs(df) <= 1e-6)) all_close = false;
                  }
                  if (dup_mode == 0 && !all_nonzero) continue;
                  if (dup_mode == 2 && (!all_nonzero || all_close)) continue;
                }
                double H[16];
                for (int c = 0; c < 4; ++c) {
                  for (int r = 0; r < 3; ++r) H[4 * c + r] = sol[12 * k + 3 * c + r];
                  H[4 * c + 3] = (c == 3) ? 1.0 : 0.0;
                }
                bool finite = true; /* isFinite (:2087-2091) */
                for (int e = 0; e < 16; ++e)
                  if (!((H[e] - H[e]) == (H[e] - H[e]))) finite = false;
                if (!finite) continue;
                st[2]++;
                double Hinv[16];
                orc_mat4_inv(H, Hinv);
                for (int m = 0; m < n_un_obj; ++m) orc_project2d(K, Hinv, obj + 4 * un_obj_idx[m], &proj[2 * m]);
                orc_min_distances_and_pairs(n_un_im, un_im.data(), n_un_obj, proj.data(), mind.data(), pr.data());
                unsigned within = 0;
                for (int ll = 0; ll < n_un_im; ++ll)
                  if (mind[ll] < tol) within++;
                if (within > 0) {
                  st[3]++;
                  for (int mm = 0; mm < 3; ++mm) hist[(size_t)ci[mm] * L + pj[mm]] += 1;
                  for (int nn = 0; nn < n_un_im; ++nn)
                    if (mind[nn] < tol) hist[(size_t)un_im_idx[nn] * L + un_obj_idx[pr[nn] - 1]] += 1;
                }
              }
            }
      }
  if (stats) std::memcpy(stats, st, sizeof st);
  return 0;
}

} /* extern "C" */

extern "C" {

/* Combinations::combinationsNoReplacement(N, K) (combinations.cpp:52-128), restated literally: rows of K 1-based
 * indices.  out: capacity rows x K; returns the number of rows. */
int orc_combinations(unsigned N, unsigned K, unsigned *out, int capacity)
{
  if (K == 1) { for (unsigned i = 0; i < N && (int)i < capacity; ++i) out[i] = i + 1; return (int)N; }
  if (K == N) { for (unsigned j = 0; j < K; ++j) out[j] = j + 1; return 1; }
  unsigned long long num = 1, den = 1;
  for (unsigned v = N - K + 1; v <= N; ++v) num *= v;   /* consecutiveProduct(N-K+1, N) */
  for (unsigned v = 1; v <= K; ++v) den *= v;           /* consecutiveProduct(1, K)     */
  const unsigned bc = (unsigned)(num / den);
  if ((int)bc > capacity) return -1;
  std::vector<unsigned> wv(K);
  for (unsigned j = 0; j < K; ++j) wv[j] = j + 1;
  unsigned lim = K, idx = 1, step, flag;
  for (unsigned j = 0; j < K; ++j) out[j] = wv[j];
  for (unsigned i = 2; i + 1 <= bc; ++i) {
    if (idx + lim < N) { step = idx; flag = 0; } else { step = 1; flag = 1; }
    for (unsigned j = 1; j <= step; ++j) wv[K + j - idx - 1] = lim + j;
    for (unsigned j = 0; j < K; ++j) out[(size_t)(i - 1) * K + j] = wv[j];
    idx = idx * flag + 1;
    lim = wv[K - idx];
  }
  for (unsigned j = 0; j < K; ++j) out[(size_t)(bc - 1) * K + j] = N - K + 1 + j;
  return (int)bc;
}

/* calculateSquaredReprojectionErrorAndCertainty (pose_estimator.cpp:1087-1132): paired points, n each */
double orc_sq_reprojection_error_and_certainty(int n, const double *image_pts, const double *object_pts, double tol,
                                               double *certainty)
{
  std::vector<double> d(n > 0 ? n : 1);
  for (int i = 0; i < n; ++i) d[i] = sqdist2(image_pts + 2 * i, object_pts + 2 * i);
  double squared_error = 0;
  unsigned num = 0;
  const double tol2 = std::pow(tol, 2);
  for (int j = 1; j <= n; ++j) {
    int r = 0;   /* minCoeff: first minimum; a NaN entry never compares smaller */
    for (int i = 1; i < n; ++i) if (d[i] < d[r]) r = i;
    const double mv = d[r];
    if (mv <= tol2) { squared_error += mv; num++; d[r] = std::numeric_limits<double>::infinity(); }
    else break;
  }
  *certainty = (double)num / n;   /* 0/0 = NaN when there is no unused correspondence: never >= a threshold */
  return squared_error;
}

/* The per-combination body of PoseEstimator::checkCorrespondences (pose_estimator.cpp:1340-1445) for one
 * correspondence set.  corr: ncorr x 2 (LED, detection), 1-based.  Per combination c (reference order):
 * valid[c] (a solution reached the certainty threshold, :1412), best[c] (its index, :1416-1420), err[c] (its squared
 * error), pose[c] (16 doubles, column-major: temp_solution.inverse() -- the particle seeded at :1437 and the
 * transform applied at :1449).  rc[c] is computePoses' return value.  Returns the number of combinations. */
int orc_check_correspondences(const double K[9], int L, const double *obj, int D, const double *img, int ncorr,
                              const unsigned *corr, double tol, double certainty_threshold, int *rc, int *valid,
                              int *best, double *err, double *pose)
{
  if (ncorr < 3) return 0;
  std::vector<double> vec(3 * (size_t)D);
  orc_image_vectors(K, D, img, vec.data());
  std::vector<unsigned> comb(3 * 1024);
  const int nc = orc_combinations((unsigned)ncorr, 3, comb.data(), 1024);
  for (int c = 0; c < nc; ++c) {
    const unsigned *cc = &comb[3 * (size_t)c];
    double fv[9], wp[9];
    for (int k = 0; k < 3; ++k) {
      const unsigned led = corr[2 * (cc[k] - 1)] - 1, det = corr[2 * (cc[k] - 1) + 1] - 1;
      for (int r = 0; r < 3; ++r) { wp[3 * k + r] = obj[4 * led + r]; fv[3 * k + r] = vec[3 * det + r]; }
    }
    double un_obj[4 * 16], un_im[2 * 16];
    int n_un = 0;
    for (int l = 0; l < ncorr; ++l) {
      if ((unsigned)l == cc[0] - 1 || (unsigned)l == cc[1] - 1 || (unsigned)l == cc[2] - 1) continue;
      const unsigned led = corr[2 * l] - 1, det = corr[2 * l + 1] - 1;
      for (int r = 0; r < 4; ++r) un_obj[4 * n_un + r] = obj[4 * led + r];
      un_im[2 * n_un] = img[2 * det]; un_im[2 * n_un + 1] = img[2 * det + 1];
      n_un++;
    }
    double sol[48];
    rc[c] = compute_poses(fv, wp, sol);
    valid[c] = 0; best[c] = -1; err[c] = std::numeric_limits<double>::infinity();
    for (int e = 0; e < 16; ++e) pose[16 * (size_t)c + e] = 0;
    if (rc[c] != 0) continue;
    for (int j = 0; j < 4; ++j) {
      double H[16];
      for (int q = 0; q < 4; ++q) {
        for (int r = 0; r < 3; ++r) H[4 * q + r] = sol[12 * j + 3 * q + r];
        H[4 * q + 3] = (q == 3) ? 1.0 : 0.0;
      }
      bool finite = true;
      for (int e = 0; e < 16; ++e) if (!((H[e] - H[e]) == (H[e] - H[e]))) finite = false;
      if (!finite) continue;
      double Hinv[16], proj[2 * 16];
      orc_mat4_inv(H, Hinv);
      for (int m = 0; m < n_un; ++m) orc_project2d(K, Hinv, un_obj + 4 * m, proj + 2 * m);
      double certainty;
      const double se = orc_sq_reprojection_error_and_certainty(n_un, un_im, proj, tol, &certainty);
      if (certainty >= certainty_threshold) {
        valid[c] = 1;
        if (se < err[c]) { err[c] = se; best[c] = j; for (int e = 0; e < 16; ++e) pose[16 * (size_t)c + e] = Hinv[e]; }
      }
    }
  }
  return nc;
}

} /* extern "C" */
