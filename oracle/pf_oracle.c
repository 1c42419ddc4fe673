/*
 * pf_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).  See pf_oracle.h.
 *
 * Every function cites the reference lines it restates
 * (/root/reference/pf_mpe_lib/src/pose_estimator.cpp unless prefixed).
 * Parity status: RNG pinned against libstdc++ (oracle/check_rng.cpp); the rest is
 * "parity unpinned" (reference has no tests/golden vectors and cannot be built here).
 *
 * Build: gcc -O3 -std=c11 -ffp-contract=off -fPIC -shared  (no FMA contraction, like the
 * reference's plain x86-64 -O3 build, pf_mpe_lib/CMakeLists.txt:5-6).
 */
#include "pf_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ======================================================================= */
/* RNG -- libstdc++ <random>, used at :476-477, :510-516, :523              */
/* ======================================================================= */
#define ORC_M 2147483647u /* 2^31-1 */
#define ORC_A 16807u

void orc_rng_seed(orc_rng *g, uint32_t seed)
{
  /* linear_congruential_engine::seed: x = seed mod m; if (c mod m == 0 && x == 0) x = 1 */
  uint32_t x = seed % ORC_M;
  if (x == 0) x = 1;
  g->x = x;
}

uint32_t orc_rng_next(orc_rng *g)
{
  g->x = (uint32_t)(((uint64_t)g->x * ORC_A) % ORC_M);
  return g->x;
}

double orc_rng_canonical(orc_rng *g)
{
  /* std::generate_canonical<double,53>(minstd_rand0): R = max-min+1 = 2147483646,
   * floor(log2 R) = 30, m = ceil(53/30) = 2 draws; sum += (g-min)*tmp; tmp *= R; ret = sum/tmp */
  const double R = 2147483646.0;
  double sum = 0.0, tmp = 1.0;
  for (int k = 0; k < 2; ++k) {
    sum += (double)(orc_rng_next(g) - 1u) * tmp;
    tmp *= R;
  }
  double ret = sum / tmp;
  if (ret >= 1.0) ret = nextafter(1.0, 0.0);
  return ret;
}

double orc_rng_uniform(orc_rng *g, double a, double b)
{
  /* uniform_real_distribution::operator(): canonical * (b - a) + a */
  return orc_rng_canonical(g) * (b - a) + a;
}

static uint32_t mulmod(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) % ORC_M); }

uint32_t orc_rng_skip(uint32_t x, uint64_t steps)
{
  uint32_t base = ORC_A, acc = 1;
  while (steps) {
    if (steps & 1) acc = mulmod(acc, base);
    base = mulmod(base, base);
    steps >>= 1;
  }
  return mulmod(x, acc);
}

/* ======================================================================= */
/* 4x4 helpers, column-major: A(r,c) = A[r + 4c]                            */
/* ======================================================================= */
#define M4(A, r, c) ((A)[(r) + 4 * (c)])

void orc_mat4_identity(double *A)
{
  memset(A, 0, 16 * sizeof(double));
  A[0] = A[5] = A[10] = A[15] = 1.0;
}

void orc_mat4_mul(const double *A, const double *B, double *C)
{
  double tmp[16];
  for (int c = 0; c < 4; ++c)
    for (int r = 0; r < 4; ++r) {
      double s = 0.0;
      for (int k = 0; k < 4; ++k) s += M4(A, r, k) * M4(B, k, c);
      tmp[r + 4 * c] = s;
    }
  memcpy(C, tmp, sizeof tmp);
}

/* general n x n inverse, Gauss-Jordan with partial pivoting; A column-major, n <= 6 */
static int mat_inv(int n, const double *A, double *Ainv)
{
  double a[36], b[36];
  for (int i = 0; i < n * n; ++i) { a[i] = A[i]; b[i] = 0.0; }
  for (int i = 0; i < n; ++i) b[i + n * i] = 1.0;
  for (int c = 0; c < n; ++c) {
    int p = c;
    double best = fabs(a[c + n * c]);
    for (int r = c + 1; r < n; ++r)
      if (fabs(a[r + n * c]) > best) { best = fabs(a[r + n * c]); p = r; }
    if (p != c)
      for (int k = 0; k < n; ++k) {
        double t = a[c + n * k]; a[c + n * k] = a[p + n * k]; a[p + n * k] = t;
        t = b[c + n * k]; b[c + n * k] = b[p + n * k]; b[p + n * k] = t;
      }
    double piv = a[c + n * c];
    double inv = 1.0 / piv; /* inf/nan propagate like an unchecked Eigen inverse */
    for (int k = 0; k < n; ++k) { a[c + n * k] *= inv; b[c + n * k] *= inv; }
    for (int r = 0; r < n; ++r) {
      if (r == c) continue;
      double f = a[r + n * c];
      if (f == 0.0) continue;
      for (int k = 0; k < n; ++k) { a[r + n * k] -= f * a[c + n * k]; b[r + n * k] -= f * b[c + n * k]; }
    }
  }
  memcpy(Ainv, b, sizeof(double) * n * n);
  return 0;
}

void orc_mat4_inv(const double *A, double *Ainv) { mat_inv(4, A, Ainv); }

/* ======================================================================= */
/* project2d  :1017-1034                                                    */
/* ======================================================================= */
void orc_project2d(const double K[9], const double T[16], const double X[4], double uv[2])
{
  /* camera_matrix = [K | 0] (3x4)  :1020-1028 ; temp = camera_matrix * transform * point
   * evaluated left to right: (K34 * T) * X   :1031 */
  double KT[12]; /* 3x4, row r col c at KT[r*4+c] */
  for (int r = 0; r < 3; ++r)
    for (int c = 0; c < 4; ++c) {
      double s = 0.0;
      for (int k = 0; k < 3; ++k) s += K[r * 3 + k] * M4(T, k, c);
      s += 0.0 * M4(T, 3, c);
      KT[r * 4 + c] = s;
    }
  double t[3];
  for (int r = 0; r < 3; ++r) {
    double s = 0.0;
    for (int c = 0; c < 4; ++c) s += KT[r * 4 + c] * X[c];
    t[r] = s;
  }
  uv[0] = t[0] / t[2]; /* :1032, no cheirality test (Q14) */
  uv[1] = t[1] / t[2];
}

/* LEDDetector::distortPoints (led_detector.cpp:371-414), one point: normalise with K, plumb_bob
 * radial (k1,k2,k3) + tangential (p1,p2) terms, back to pixels.  D5 = (k1,k2,p1,p2,k3). */
void orc_distort_point(const double K[9], const double D5[5], const double in[2], double out[2])
{
  const double fx = K[0], fy = K[4], cx = K[2], cy = K[5];
  const double k1 = D5[0], k2 = D5[1], p1 = D5[2], p2 = D5[3], k3 = D5[4];
  const double x = (in[0] - cx) / fx, y = (in[1] - cy) / fy;
  const double r2 = x * x + y * y;
  double xc = x * (1. + k1 * r2 + k2 * r2 * r2 + k3 * r2 * r2 * r2);
  double yc = y * (1. + k1 * r2 + k2 * r2 * r2 + k3 * r2 * r2 * r2);
  xc = xc + (2. * p1 * x * y + p2 * (r2 + 2. * x * x));
  yc = yc + (p1 * (r2 + 2. * y * y) + 2. * p2 * x * y);
  out[0] = xc * fx + cx;
  out[1] = yc * fy + cy;
}

/* ======================================================================= */
/* calculateEstimationProbability  :2385-2445                               */
/* ======================================================================= */
double orc_score(const orc_params *prm, int L, int D, const double *img, const double *proj,
                 unsigned *pairs, int *npairs)
{
  double P = 0.0;
  int ncorr = 0;
  if (D <= 0 || L <= 0) { *npairs = 0; return 0.0; }
  /* distances(i,j), i = detection (row), j = LED (col), column-major  :2393-2401 */
  double *dist = (double *)malloc(sizeof(double) * (size_t)D * (size_t)L);
  for (int i = 0; i < D; ++i)
    for (int j = 0; j < L; ++j) {
      double dx = img[2 * i] - proj[2 * j], dy = img[2 * i + 1] - proj[2 * j + 1];
      dist[i + (size_t)D * j] = dx * dx + dy * dy; /* (p1-p2).squaredNorm()  :2084 */
    }
  unsigned used[ORC_MAX_LED];
  int nused = 0, self_occ = 1;
  int steps = D < L ? D : L; /* :2409 */
  for (int k = 0; k < steps; ++k) {
    /* Eigen minCoeff(&row,&col): column-major visit, strict '<' -> first minimum  :2411 */
    double mv = dist[0];
    int ri = 0, ci = 0;
    for (int j = 0; j < L; ++j)
      for (int i = 0; i < D; ++i) {
        double v = dist[i + (size_t)D * j];
        if (v < mv) { mv = v; ri = i; ci = j; }
      }
    mv = sqrt(mv); /* :2412 */
    if (mv <= prm->back_projection_pixel_tolerance_PF) { /* :2414 */
      double g = (prm->back_projection_pixel_tolerance - mv) / prm->back_projection_pixel_tolerance;
      P += (double)L + g * g; /* pow(.,2)  :2416 ; object_points_.size() == L */
      pairs[2 * k] = (unsigned)ci + 1u;     /* :2417 */
      pairs[2 * k + 1] = (unsigned)ri + 1u; /* :2418 */
      ++ncorr;
      int reused = 0;
      for (int u = 0; u < nused; ++u) if (used[u] == (unsigned)ri) reused = 1; /* :2421 */
      if (reused) { P = P - self_occ * 3; self_occ++; } /* :2423-2424 */
      used[nused++] = (unsigned)ri; /* :2427 */
      if (ci < 5 && prm->marker_downgrade[ci]) P = P - 2; /* :2431-2432 (vector has exactly 5 entries, Q5) */
      for (int i = 0; i < D; ++i) dist[i + (size_t)D * ci] = INFINITY; /* :2435 */
    } else
      break; /* :2438 */
  }
  free(dist);
  *npairs = ncorr; /* conservativeResize  :2441 */
  return P;
}

/* ======================================================================= */
/* stratified resampling  :668-682                                          */
/* ======================================================================= */
void orc_resample(int N, const double *w, const double *r, int literal, int *idx, unsigned *count)
{
  for (int i = 0; i < N; ++i) count[i] = 0;
  int particle_index = -1; /* reference: uninitialised `int Particle_index` (:517) */
  if (literal) {
    for (int k = 0; k < N; ++k) {
      double u = ((double)k + r[k]) / (double)N; /* :671 */
      double c = 0.0;                            /* :672 */
      for (int i = 0; i < N; ++i) {              /* :674-679 */
        c += w[i];
        if (c >= u) { particle_index = i; count[i]++; break; }
      }
      idx[k] = particle_index; /* :681 */
    }
  } else {
    /* O(N) equivalent: the sequential prefix c_i is the same for every k, and u_k is
     * non-decreasing in k, so "first i with c_i >= u_k" is a first crossing of the running
     * maximum of the prefix; a single forward sweep over (i) serves all k in order. */
    double *c = (double *)malloc(sizeof(double) * (size_t)N);
    double acc = 0.0;
    for (int i = 0; i < N; ++i) { acc += w[i]; c[i] = acc; }
    int i = 0;
    double prev_u = -INFINITY;
    for (int k = 0; k < N; ++k) {
      double u = ((double)k + r[k]) / (double)N;
      if (u < prev_u) i = 0; /* cannot happen for r in [0,1); keeps the sweep exact if it did */
      prev_u = u;
      while (i < N && !(c[i] >= u)) ++i;
      if (i < N) { particle_index = i; count[i]++; }
      idx[k] = particle_index;
    }
    free(c);
  }
}

/* ======================================================================= */
/* the PF block  :475-733                                                   */
/* ======================================================================= */
int orc_pf_update(const orc_params *prm, const orc_pf_in *in, orc_pf_out *out)
{
  const int N = in->N, L = in->L, D = in->D;
  if (N <= 0 || L <= 0 || L > ORC_MAX_LED) return -1;
  orc_rng g;
  orc_rng_seed(&g, in->seed); /* :476-477 with injected seed */
  uint64_t steps0 = 0;

  /* range factors  :488-506 (Q1: y and z reuse predictionMatrix(0,3)) */
  double facTX, facTY, facTZ, facRX, facRY, facRZ;
  if (in->isi == 1) {
    facTX = facTY = facTZ = 1; facRX = facRY = facRZ = 1;
  } else {
    double timeDiffFrames = in->dt;
    double f = fmin(fmax(0.2, fabs(M4(in->P, 0, 3)) / timeDiffFrames), 1.0) / 4;
    facTX = f; facTY = f; facTZ = f;
    facRX = facRY = facRZ = 0.2;
  }
  /* distributions  :510-516 */
  const double tXa = prm->min_transition_noise * facTX, tXb = prm->max_transition_noise * facTX;
  const double tYa = prm->min_transition_noise * facTY, tYb = prm->max_transition_noise * facTY;
  const double tZa = prm->min_transition_noise * facTZ, tZb = prm->max_transition_noise * facTZ;
  const double aXa = prm->min_angular_noise * facRX, aXb = prm->max_angular_noise * facRX;
  const double aYa = prm->min_angular_noise * facRY, aYb = prm->max_angular_noise * facRY;
  const double aZa = prm->min_angular_noise * facRZ, aZb = prm->max_angular_noise * facRZ;

  double highestProb = 0; /* :519 */
  int iter = 0;           /* :527 */
  int mostLikelyParticleIdx = -1;
  int have_likely = 0, best_iter = -1;
  const size_t PS = 2 * ORC_MAX_LED;

  double *T = (double *)malloc(sizeof(double) * 16 * (size_t)N);       /* PoseParticle */
  double *w = (double *)calloc((size_t)N, sizeof(double));             /* probPart (:520 setZero) */
  unsigned *pairs = (unsigned *)calloc(PS * (size_t)N, sizeof(unsigned));
  int *npairs = (int *)calloc((size_t)N, sizeof(int));
  double proj[2 * ORC_MAX_LED];
  double maxw;
  const int max_iter = in->max_iter > 0 ? in->max_iter : 80;
  int numLED = D; /* numLED = detected_led_positions.size()  :449 */

  do { /* :535 */
    for (int n = 0; n < N; ++n) { /* :543 */
      double *Tn = T + 16 * (size_t)n;
      if (n == 0) {
        memcpy(Tn, in->T_cur, 16 * sizeof(double)); /* :547 */
      } else if (n == 1) {
        memcpy(Tn, in->T_pred, 16 * sizeof(double)); /* :551 */
      } else {
        double B[16], tmp[16];
        const double *En = in->E + 16 * (size_t)n;
        if (in->isi > 1 && (iter % 10) != 0) { /* :555-556, left-to-right */
          orc_mat4_mul(in->C, En, tmp);
          orc_mat4_mul(tmp, in->P, B);
        } else if (in->isi > 1) { /* :557-558 */
          orc_mat4_mul(in->C, En, B);
        } else {
          memcpy(B, En, sizeof B); /* :560 */
        }
        const double s = 1 + 0.025 * floor((double)(iter / 10)); /* integer division then floor */
        double a = orc_rng_uniform(&g, aXa, aXb) * s; /* :563 */
        double b = orc_rng_uniform(&g, aYa, aYb) * s; /* :564 */
        double c = orc_rng_uniform(&g, aZa, aZb) * s; /* :565 */
        steps0 += 6;
        double rotX[16], rotY[16], rotZ[16];
        orc_mat4_identity(rotX); orc_mat4_identity(rotY); orc_mat4_identity(rotZ);
        M4(rotX, 1, 1) = cos(a); M4(rotX, 1, 2) = -sin(a); M4(rotX, 2, 1) = sin(a); M4(rotX, 2, 2) = cos(a); /* :567-570 */
        M4(rotY, 0, 0) = cos(b); M4(rotY, 0, 2) = sin(b); M4(rotY, 2, 0) = -sin(b); M4(rotY, 2, 2) = cos(b); /* :572-575 */
        M4(rotZ, 0, 0) = cos(c); M4(rotZ, 0, 1) = -sin(c); M4(rotZ, 1, 0) = sin(c); M4(rotZ, 1, 1) = cos(c); /* :577-580 */
        orc_mat4_mul(B, rotZ, tmp);   /* :582 ((B*rotZ)*rotY)*rotX */
        orc_mat4_mul(tmp, rotY, tmp);
        orc_mat4_mul(tmp, rotX, Tn);
        M4(Tn, 0, 3) = M4(B, 0, 3) + orc_rng_uniform(&g, tXa, tXb) * s; /* :585 */
        M4(Tn, 1, 3) = M4(B, 1, 3) + orc_rng_uniform(&g, tYa, tYb) * s; /* :586 */
        M4(Tn, 2, 3) = M4(B, 2, 3) + orc_rng_uniform(&g, tZa, tZb) * s; /* :587 */
        steps0 += 6;
      }
      for (int j = 0; j < L; ++j) {
        orc_project2d(in->K, Tn, in->obj + 4 * j, proj + 2 * j); /* :596-599 */
        if (in->dist) orc_distort_point(in->K, in->dist, proj + 2 * j, proj + 2 * j); /* option f3, not in the reference PF */
      }
      w[n] = orc_score(prm, L, D, in->img, proj, pairs + PS * (size_t)n, npairs + n);      /* :602 */
    }
    iter++; /* :606 */
    maxw = w[0];
    int arg = 0;
    for (int n = 1; n < N; ++n) if (w[n] > maxw) { maxw = w[n]; arg = n; } /* maxCoeff: first max */
    if (maxw > highestProb) { /* :608-614 */
      highestProb = maxw;
      mostLikelyParticleIdx = arg;
      memcpy(out->T, T, sizeof(double) * 16 * (size_t)N);
      memcpy(out->w_raw, w, sizeof(double) * (size_t)N);
      memcpy(out->pairs, pairs, sizeof(unsigned) * PS * (size_t)N);
      memcpy(out->npairs, npairs, sizeof(int) * (size_t)N);
      have_likely = 1;
      best_iter = iter - 1;
    }
  } while (iter < max_iter && maxw < (double)((long)L * (long)(5 < numLED ? 5 : numLED))); /* :616 */

  if (!have_likely) { /* :619-624: nothing to restore -> last iteration's arrays stay */
    memcpy(out->T, T, sizeof(double) * 16 * (size_t)N);
    memcpy(out->w_raw, w, sizeof(double) * (size_t)N);
    memcpy(out->pairs, pairs, sizeof(unsigned) * PS * (size_t)N);
    memcpy(out->npairs, npairs, sizeof(int) * (size_t)N);
  }
  free(T); free(w); free(pairs); free(npairs);

  /* :627-629 */
  double sum = 0.0;
  for (int n = 0; n < N; ++n) sum += out->w_raw[n];
  if (sum != 0) for (int n = 0; n < N; ++n) out->w[n] = out->w_raw[n] / sum;
  else          for (int n = 0; n < N; ++n) out->w[n] = out->w_raw[n];

  out->iterations = iter;
  out->best_iteration = best_iter;
  out->highest_prob = highestProb;
  out->weight_sum = sum;
  out->most_likely_particle_idx = mostLikelyParticleIdx;
  out->most_likely_pose_idx = -1;
  out->resampled = 0;
  out->winner_npairs = 0;
  out->isi_out = in->isi;
  for (int n = 0; n < N; ++n) { out->idx[n] = -1; out->count[n] = 0; }

  /* :633 */
  if (sum != 0 && highestProb > (double)((long)L * (long)(3 < numLED ? 3 : numLED))) {
    out->accepted = 1;
    out->flag_fail = 1.0f; /* :635; branch :637-661 is dead code (2/3*numLED == 0, Q2) */
    if (in->isi > 0) {     /* :666 */
      double *r = (double *)malloc(sizeof(double) * (size_t)N);
      for (int k = 0; k < N; ++k) r[k] = orc_rng_uniform(&g, 0.0, 1.0); /* :671 draws, in k order */
      steps0 += 2 * (uint64_t)N;
      orc_resample(N, out->w, r, in->resampler == 0, out->idx, out->count);
      free(r);
      /* :681 newPoseEstimation[k] = PoseParticle[Particle_index]; idx -1 (undefined in the
       * reference) is mapped to "leave E_new[k] as the input E[k]" */
      for (int k = 0; k < N; ++k) {
        const double *src = out->idx[k] >= 0 ? out->T + 16 * (size_t)out->idx[k] : in->E + 16 * (size_t)k;
        memcpy(out->E_new + 16 * (size_t)k, src, 16 * sizeof(double));
      }
      /* :685-688 first max of counterMeas */
      unsigned mc = out->count[0];
      int q = 0;
      for (int i = 1; i < N; ++i) if (out->count[i] > mc) { mc = out->count[i]; q = i; }
      out->most_likely_pose_idx = q;
      memcpy(out->winner_pose, out->T + 16 * (size_t)q, 16 * sizeof(double));
      out->winner_npairs = out->npairs[q];
      memcpy(out->winner_pairs, out->pairs + PS * (size_t)q, sizeof(unsigned) * PS);
      out->resampled = 1;
    }
  } else { /* :707-719 */
    out->accepted = 0;
    out->flag_fail = 4.0f;
    out->isi_out = 0;
    double mw = out->w[0];
    int arg = 0;
    for (int n = 1; n < N; ++n) if (out->w[n] > mw) { mw = out->w[n]; arg = n; } /* :714 on normalised probPart */
    out->most_likely_particle_idx = arg;
    memcpy(out->winner_pose, out->T + 16 * (size_t)arg, 16 * sizeof(double)); /* :716 */
    if (in->E != out->E_new) memcpy(out->E_new, in->E, sizeof(double) * 16 * (size_t)N);
  }
  out->rng_steps = steps0;
  return 0;
}

/* ======================================================================= */
/* Lie maps  :2194-2303                                                     */
/* ======================================================================= */
static void skew(const double w[3], double O[9]) /* 3x3 col-major  :2298-2303 */
{
  /* Omega << 0, -w2, w1,  w2, 0, -w0,  -w1, w0, 0 (row-wise) */
  O[0] = 0;     O[3] = -w[2]; O[6] = w[1];
  O[1] = w[2];  O[4] = 0;     O[7] = -w[0];
  O[2] = -w[1]; O[5] = w[0];  O[8] = 0;
}
static void m3mul(const double *A, const double *B, double *C)
{
  double t[9];
  for (int c = 0; c < 3; ++c)
    for (int r = 0; r < 3; ++r) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += A[r + 3 * k] * B[k + 3 * c];
      t[r + 3 * c] = s;
    }
  memcpy(C, t, sizeof t);
}

void orc_exponential_map(const double twist[6], double T[16])
{
  const double *ups = twist, *om = twist + 3;
  double theta = sqrt(om[0] * om[0] + om[1] * om[1] + om[2] * om[2]);
  double th2 = theta * theta;
  double O[9], O2[9], R[9], V[9];
  skew(om, O);
  m3mul(O, O, O2);
  for (int i = 0; i < 9; ++i) { R[i] = 0; V[i] = 0; }
  R[0] = R[4] = R[8] = 1; V[0] = V[4] = V[8] = 1;
  if (theta != 0) { /* :2212-2218 */
    for (int i = 0; i < 9; ++i) {
      R[i] += O[i] / theta * sin(theta) + O2[i] / th2 * (1 - cos(theta));
      V[i] += (1 - cos(theta)) / th2 * O[i] + (theta - sin(theta)) / (th2 * theta) * O2[i];
    }
  }
  orc_mat4_identity(T);
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) M4(T, r, c) = R[r + 3 * c];
  for (int r = 0; r < 3; ++r) M4(T, r, 3) = V[r] * ups[0] + V[r + 3] * ups[1] + V[r + 6] * ups[2];
}

void orc_logarithm_map(const double T[16], double xi[6])
{
  double R[9], t[3], w_hat[9], A_inv[9];
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) R[r + 3 * c] = M4(T, r, c);
  for (int r = 0; r < 3; ++r) t[r] = M4(T, r, 3);
  double phi = 0;
  /* R.isApprox(I, 1e-10): ||R-I||_F^2 <= p^2 * min(||R||_F^2, ||I||_F^2)   :2240 */
  double d2 = 0, r2 = 0;
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) {
    double e = R[r + 3 * c] - (r == c ? 1.0 : 0.0);
    d2 += e * e; r2 += R[r + 3 * c] * R[r + 3 * c];
  }
  int approxI = d2 <= 1e-10 * 1e-10 * fmin(r2, 3.0);
  for (int i = 0; i < 9; ++i) w_hat[i] = 0;
  if (!approxI) {
    double temp = (R[0] + R[4] + R[8] - 1) / 2; /* :2247 */
    if (temp > 1) temp = 1; else if (temp < -1) temp = -1;
    phi = acos(temp);
    if (phi != 0)
      for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r)
        w_hat[r + 3 * c] = (R[r + 3 * c] - R[c + 3 * r]) / (2 * sin(phi)) * phi; /* :2265 */
  }
  double w[3] = { w_hat[2 + 3 * 1], w_hat[0 + 3 * 2], w_hat[1 + 3 * 0] }; /* :2270 */
  double w_norm = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  /* t.isApproxToConstant(0, 1e-10): isApprox(t, zeros): ||t||^2 <= p^2 * min(||t||^2, 0) -> only exact zero */
  double tn2 = t[0] * t[0] + t[1] * t[1] + t[2] * t[2];
  if (tn2 <= 1e-20 * fmin(tn2, 0.0)) {
    for (int i = 0; i < 9; ++i) A_inv[i] = 0; /* :2277 */
  } else if (w_norm == 0 || sin(w_norm) == 0) {
    for (int i = 0; i < 9; ++i) A_inv[i] = 0;
    A_inv[0] = A_inv[4] = A_inv[8] = 1; /* :2281 */
  } else {
    double wh2[9];
    m3mul(w_hat, w_hat, wh2);
    double f = (2 * sin(w_norm) - w_norm * (1 + cos(w_norm))) / (2 * w_norm * w_norm * sin(w_norm));
    for (int i = 0; i < 9; ++i) A_inv[i] = -w_hat[i] / 2 + f * wh2[i];
    A_inv[0] += 1; A_inv[4] += 1; A_inv[8] += 1; /* :2285-2286 */
  }
  for (int r = 0; r < 3; ++r) xi[r] = A_inv[r] * t[0] + A_inv[r + 3] * t[1] + A_inv[r + 6] * t[2];
  xi[3] = w[0]; xi[4] = w[1]; xi[5] = w[2];
}

void orc_predict_pose(const double prev[16], const double cur[16], double t_prev, double t_cur,
                      double t_pred, double P[16], double pred[16])
{
  /* :1000-1007 */
  double inv[16], rel[16], delta[6], dh[6];
  orc_mat4_inv(prev, inv);
  orc_mat4_mul(inv, cur, rel);
  orc_logarithm_map(rel, delta);
  for (int i = 0; i < 6; ++i) dh[i] = delta[i] / (t_cur - t_prev) * (t_pred - t_cur);
  orc_exponential_map(dh, P);
  orc_mat4_mul(cur, P, pred);
}

/* ======================================================================= */
/* Gauss-Newton  :1805-2009, computeJacobian :2163-2192                     */
/* ======================================================================= */
static void ldlt_solve6(const double *A, const double *b, double *x)
{
  /* Eigen's A.ldlt().solve(b): LDL^T with symmetric (diagonal) pivoting. */
  const int n = 6;
  double a[36];
  int perm[6];
  memcpy(a, A, sizeof a);
  for (int i = 0; i < n; ++i) perm[i] = i;
  for (int k = 0; k < n; ++k) {
    int p = k;
    double best = fabs(a[k + n * k]);
    for (int i = k + 1; i < n; ++i) if (fabs(a[i + n * i]) > best) { best = fabs(a[i + n * i]); p = i; }
    if (p != k) {
      for (int j = 0; j < n; ++j) { double t = a[k + n * j]; a[k + n * j] = a[p + n * j]; a[p + n * j] = t; }
      for (int j = 0; j < n; ++j) { double t = a[j + n * k]; a[j + n * k] = a[j + n * p]; a[j + n * p] = t; }
      int t = perm[k]; perm[k] = perm[p]; perm[p] = t;
    }
    double d = a[k + n * k];
    if (d == 0.0) continue;
    for (int i = k + 1; i < n; ++i) a[i + n * k] /= d;
    for (int j = k + 1; j < n; ++j)
      for (int i = j; i < n; ++i) {
        a[i + n * j] -= a[i + n * k] * d * a[j + n * k];
        a[j + n * i] = a[i + n * j];
      }
  }
  double y[6];
  for (int i = 0; i < n; ++i) y[i] = b[perm[i]];
  for (int i = 0; i < n; ++i) for (int k = 0; k < i; ++k) y[i] -= a[i + n * k] * y[k];
  for (int i = 0; i < n; ++i) y[i] = (a[i + n * i] != 0.0) ? y[i] / a[i + n * i] : 0.0;
  for (int i = n - 1; i >= 0; --i) for (int k = i + 1; k < n; ++k) y[i] -= a[k + n * i] * y[k];
  for (int i = 0; i < n; ++i) x[perm[i]] = y[i];
}

int orc_optimise_pose(const double K[9], int L, const double *obj, int D, const double *img,
                      const unsigned *pairs, int npairs, double pose[16], double cov[36])
{
  (void)L; (void)D;
  const double converged = 1e-13;
  const unsigned max_itr = 500;
  double pose_init[16];
  memcpy(pose_init, pose, sizeof pose_init);
  const double fx = K[0], fy = K[4]; /* focal_lengths  :1821-1822 */
  double e_init = 0, e_end = 0;
  double A[36], b[6], dT[6];
  int num_iter = -1;
  for (int i = 0; i < 36; ++i) A[i] = 0;
  for (unsigned i = 0; i < max_itr; ++i) {
    for (int k = 0; k < 36; ++k) A[k] = 0;
    for (int k = 0; k < 6; ++k) b[k] = 0;
    for (int j = 0; j < npairs; ++j) {
      if (pairs[2 * j + 1] == 0) continue; /* :1847 */
      const double *X = obj + 4 * (pairs[2 * j] - 1);
      double uv[2];
      orc_project2d(K, pose, X, uv); /* :1851 */
      double e[2] = { img[2 * (pairs[2 * j + 1] - 1)] - uv[0], img[2 * (pairs[2 * j + 1] - 1) + 1] - uv[1] };
      double en = sqrt(e[0] * e[0] + e[1] * e[1]);
      if (i == 0) e_init = +en;                 /* `=+` sic  :1859 */
      else if (i + 1 == max_itr) e_end = +en;   /* :1861 */
      /* computeJacobian  :2170-2188 */
      double pc[4];
      for (int r = 0; r < 4; ++r) pc[r] = M4(pose, r, 0) * X[0] + M4(pose, r, 1) * X[1] + M4(pose, r, 2) * X[2] + M4(pose, r, 3) * X[3];
      double x = pc[0], y = pc[1], z = pc[2], z2 = z * z;
      double J[12]; /* 2x6 row-major */
      J[0] = 1 / z * fx; J[1] = 0; J[2] = -x / z2 * fx; J[3] = -x * y / z2 * fx; J[4] = (1 + (x * x / z2)) * fx; J[5] = -y / z * fx;
      J[6] = 0; J[7] = 1 / z * fy; J[8] = -y / z2 * fy; J[9] = -(1 + y * y / z2) * fy; J[10] = x * y / z2 * fy; J[11] = x / z * fy;
      /* A += J^T R^-1 J, b += J^T R^-1 e with R = I  :1868-1869 */
      for (int c = 0; c < 6; ++c)
        for (int r = 0; r < 6; ++r) A[r + 6 * c] += J[r] * J[c] + J[6 + r] * J[6 + c];
      for (int r = 0; r < 6; ++r) b[r] += J[r] * e[0] + J[6 + r] * e[1];
    }
    ldlt_solve6(A, b, dT); /* :1872 */
    double ex[16], Tn[16];
    orc_exponential_map(dT, ex);
    orc_mat4_mul(ex, pose, Tn); /* :1876 */
    memcpy(pose, Tn, sizeof Tn);
    double nm = -1;
    for (int k = 0; k < 6; ++k) { double v = fabs(dT[k]); if (v > nm) nm = v; } /* norm_max :2305 */
    if (nm <= converged) { num_iter = (int)i; break; } /* :1881-1885 (NaN never converges) */
    if (i + 1 == max_itr) {
      if (e_init < e_end) memcpy(pose, pose_init, sizeof pose_init); /* :1888-1892 */
    }
  }
  mat_inv(6, A, cov); /* pose_covariance_ = A.inverse()  :2004 */
  return num_iter;
}

/* ======================================================================= */
/* tracked frame: estimateBodyPose PF branch  :223-733 w/o detection / ROI  */
/* ======================================================================= */
int orc_tracker_init(orc_tracker *t, int N, int L, const double *obj, const double K[9],
                     const orc_params *prm, const double pose0[16], double t0,
                     const double pose1[16], double t1, int resampler)
{
  memset(t, 0, sizeof *t);
  if (L > ORC_MAX_LED) return -1;
  t->N = N; t->L = L;
  memcpy(t->obj, obj, sizeof(double) * 4 * (size_t)L);
  memcpy(t->K, K, sizeof(double) * 9);
  t->prm = *prm;
  memcpy(t->previous_pose, pose0, sizeof t->previous_pose);
  memcpy(t->current_pose, pose1, sizeof t->current_pose);
  memcpy(t->predicted_pose, pose1, sizeof t->predicted_pose);
  t->previous_time = t0; t->current_time = t1; t->predicted_time = t1;
  t->isi = 2;
  t->resampler = resampler;
  size_t n = (size_t)N;
  t->E = (double *)malloc(sizeof(double) * 16 * n);
  t->T = (double *)malloc(sizeof(double) * 16 * n);
  t->w = (double *)malloc(sizeof(double) * n);
  for (size_t i = 0; i < n; ++i) { memcpy(t->E + 16 * i, pose1, 16 * sizeof(double)); memcpy(t->T + 16 * i, pose1, 16 * sizeof(double)); }
  orc_pf_out *o = &t->pf;
  o->T = t->T; o->w = t->w;
  o->w_raw = (double *)malloc(sizeof(double) * n);
  o->pairs = (unsigned *)malloc(sizeof(unsigned) * 2 * ORC_MAX_LED * n);
  o->npairs = (int *)malloc(sizeof(int) * n);
  o->E_new = (double *)malloc(sizeof(double) * 16 * n);
  o->idx = (int *)malloc(sizeof(int) * n);
  o->count = (unsigned *)malloc(sizeof(unsigned) * n);
  return 0;
}

void orc_tracker_free(orc_tracker *t)
{
  free(t->E); free(t->T); free(t->w);
  free(t->pf.w_raw); free(t->pf.pairs); free(t->pf.npairs); free(t->pf.E_new); free(t->pf.idx); free(t->pf.count);
  memset(t, 0, sizeof *t);
}

float orc_tracker_frame(orc_tracker *t, double time_to_predict, int D, const double *img, uint32_t seed)
{
  double P[16], C[16];
  /* predictPose  :234 -> :995-1010 */
  t->predicted_time = time_to_predict;
  orc_predict_pose(t->previous_pose, t->current_pose, t->previous_time, t->current_time, t->predicted_time, P, t->predicted_pose);
  orc_mat4_identity(C); /* bUseCamPos == false  :393 */
  orc_mat4_mul(C, t->predicted_pose, t->predicted_pose); /* :395 */

  orc_pf_in in;
  memset(&in, 0, sizeof in);
  in.N = t->N; in.L = t->L; in.D = D; in.obj = t->obj; in.img = img; in.K = t->K; in.E = t->E;
  in.T_cur = t->current_pose; in.T_pred = t->predicted_pose; in.P = P; in.C = C;
  in.isi = t->isi; in.dt = t->predicted_time - t->current_time; in.seed = seed;
  in.resampler = t->resampler; in.max_iter = 80;
  orc_pf_update(&t->prm, &in, &t->pf);
  float flag = t->pf.flag_fail;
  t->isi = t->pf.isi_out;
  if (t->pf.resampled) {
    memcpy(t->E, t->pf.E_new, sizeof(double) * 16 * (size_t)t->N);
    double before[16];
    memcpy(t->predicted_pose, t->pf.winner_pose, sizeof before); /* :687 */
    memcpy(before, t->predicted_pose, sizeof before);
    /* optimiseAndUpdatePose  :690 -> :2023-2035 */
    orc_optimise_pose(t->K, t->L, t->obj, D, img, t->pf.winner_pairs, t->pf.winner_npairs, t->predicted_pose, t->pose_covariance);
    if (t->isi < 2) t->isi++;
    /* updatePose  :2011-2021 */
    memcpy(t->previous_pose, t->current_pose, sizeof before);
    memcpy(t->current_pose, t->predicted_pose, sizeof before);
    if (t->predicted_time - t->current_time > 0.001 || t->predicted_time < t->current_time) {
      t->previous_time = t->current_time;
      t->current_time = t->predicted_time;
    }
    /* jump check  :693-701 */
    for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c)
      if (fabs(M4(before, r, c) - M4(t->predicted_pose, r, c)) >= 0.3) flag = 1.5f;
  } else if (!t->pf.accepted) {
    memcpy(t->predicted_pose, t->pf.winner_pose, 16 * sizeof(double)); /* :716 */
  }
  return flag;
}

/* ======================================================================= */
/* weighted mean pose + covariance -- north_star extra         */
/* ======================================================================= */
static void jacobi_eig3(const double S[9], double V[9], double ev[3])
{
  /* symmetric 3x3 eigen-decomposition (cyclic Jacobi), column-major */
  double a[9];
  memcpy(a, S, sizeof a);
  for (int i = 0; i < 9; ++i) V[i] = 0;
  V[0] = V[4] = V[8] = 1;
  for (int sweep = 0; sweep < 64; ++sweep) {
    double off = a[3] * a[3] + a[6] * a[6] + a[7] * a[7];
    if (off < 1e-300) break;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        double apq = a[p + 3 * q];
        if (apq == 0) continue;
        double th = (a[q + 3 * q] - a[p + 3 * p]) / (2 * apq);
        double tt = (th >= 0 ? 1.0 : -1.0) / (fabs(th) + sqrt(th * th + 1));
        double c = 1 / sqrt(tt * tt + 1), s = tt * c;
        for (int k = 0; k < 3; ++k) { /* columns */
          double akp = a[k + 3 * p], akq = a[k + 3 * q];
          a[k + 3 * p] = c * akp - s * akq; a[k + 3 * q] = s * akp + c * akq;
        }
        for (int k = 0; k < 3; ++k) { /* rows */
          double apk = a[p + 3 * k], aqk = a[q + 3 * k];
          a[p + 3 * k] = c * apk - s * aqk; a[q + 3 * k] = s * apk + c * aqk;
        }
        for (int k = 0; k < 3; ++k) {
          double vkp = V[k + 3 * p], vkq = V[k + 3 * q];
          V[k + 3 * p] = c * vkp - s * vkq; V[k + 3 * q] = s * vkp + c * vkq;
        }
      }
  }
  ev[0] = a[0]; ev[1] = a[4]; ev[2] = a[8];
}

static void project_so3(const double M[9], double R[9])
{
  /* nearest rotation to M (col-major): R = U diag(1,1,det(UV^T)) V^T via eig of M^T M */
  double MtM[9], V[9], ev[3];
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) {
    double s = 0;
    for (int k = 0; k < 3; ++k) s += M[k + 3 * r] * M[k + 3 * c];
    MtM[r + 3 * c] = s;
  }
  jacobi_eig3(MtM, V, ev);
  /* sort descending */
  for (int i = 0; i < 2; ++i) for (int j = i + 1; j < 3; ++j) if (ev[j] > ev[i]) {
    double t = ev[i]; ev[i] = ev[j]; ev[j] = t;
    for (int k = 0; k < 3; ++k) { double v = V[k + 3 * i]; V[k + 3 * i] = V[k + 3 * j]; V[k + 3 * j] = v; }
  }
  double detV = V[0] * (V[4] * V[8] - V[7] * V[5]) - V[3] * (V[1] * V[8] - V[7] * V[2]) + V[6] * (V[1] * V[5] - V[4] * V[2]);
  if (detV < 0) for (int k = 0; k < 3; ++k) V[k + 6] = -V[k + 6];
  double U[9];
  for (int j = 0; j < 2; ++j) {
    double s = sqrt(ev[j] > 0 ? ev[j] : 0);
    for (int r = 0; r < 3; ++r) {
      double v = 0;
      for (int k = 0; k < 3; ++k) v += M[r + 3 * k] * V[k + 3 * j];
      U[r + 3 * j] = s > 0 ? v / s : 0;
    }
  }
  /* third column: u0 x u1 gives det(U) = +1; sign of sigma_3 = sign(det M) handled below */
  U[6] = U[1] * U[5] - U[2] * U[4];
  U[7] = U[2] * U[3] - U[0] * U[5];
  U[8] = U[0] * U[4] - U[1] * U[3];
  /* with det(V) = +1 and det(U) = +1, R = U V^T is the det=+1 closest rotation */
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) {
    double s = 0;
    for (int k = 0; k < 3; ++k) s += U[r + 3 * k] * V[c + 3 * k];
    R[r + 3 * c] = s;
  }
}

void orc_weighted_mean_cov(int N, const double *T, const double *w, double mean_pose[16], double cov[36])
{
  double tb[3] = { 0, 0, 0 }, M[9], Rb[9];
  for (int i = 0; i < 9; ++i) M[i] = 0;
  for (int n = 0; n < N; ++n) {
    const double *Tn = T + 16 * (size_t)n;
    for (int r = 0; r < 3; ++r) tb[r] += w[n] * M4(Tn, r, 3);
    for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) M[r + 3 * c] += w[n] * M4(Tn, r, c);
  }
  project_so3(M, Rb);
  orc_mat4_identity(mean_pose);
  for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) M4(mean_pose, r, c) = Rb[r + 3 * c];
  for (int r = 0; r < 3; ++r) M4(mean_pose, r, 3) = tb[r];
  for (int i = 0; i < 36; ++i) cov[i] = 0;
  for (int n = 0; n < N; ++n) {
    const double *Tn = T + 16 * (size_t)n;
    double D4[16], xi6[6], xi[6];
    orc_mat4_identity(D4);
    for (int c = 0; c < 3; ++c) for (int r = 0; r < 3; ++r) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += Rb[k + 3 * r] * M4(Tn, k, c); /* Rbar^T R_n */
      M4(D4, r, c) = s;
    }
    orc_logarithm_map(D4, xi6); /* translation part zero -> only omega used */
    for (int r = 0; r < 3; ++r) xi[r] = M4(Tn, r, 3) - tb[r];
    xi[3] = xi6[3]; xi[4] = xi6[4]; xi[5] = xi6[5];
    for (int c = 0; c < 6; ++c) for (int r = 0; r < 6; ++r) cov[r + 6 * c] += w[n] * xi[r] * xi[c];
  }
}

/* ======================================================================= */
/* ROI prediction                                     */
/* ======================================================================= */
/* predictMarkerPositionsInImage (pose_estimator.cpp:1036-1052): every LED through every resampled
 * particle as camMoveInv * newPoseEstimation[j] * predictionMatrix, then through predicted_pose_;
 * followed by the extreme-value loop of LEDDetector::determineROI (led_detector.cpp:221-244):
 * x_min, y_min start at +INF, x_max, y_max start at 0, strict comparisons (a NaN never wins).
 * out = {x_min, y_min, x_max, y_max}: the two corner points determineROI goes on to distort/clamp. */
void orc_predict_roi(int N, const double *E, const double Cm[16], const double P[16], const double pred[16],
                     const double *markers_xyz, int L, const double K[9], double out[4])
{
  double x_min = INFINITY, x_max = 0, y_min = INFINITY, y_max = 0;
  for (int i = 0; i < L; ++i) {
    const double X[4] = { markers_xyz[3 * i], markers_xyz[3 * i + 1], markers_xyz[3 * i + 2], 1.0 };
    for (int j = 0; j <= N; ++j) {
      double T[16], uv[2];
      if (j < N) {
        double CE[16];
        orc_mat4_mul(Cm, E + 16 * (size_t)j, CE);
        orc_mat4_mul(CE, P, T);
      } else memcpy(T, pred, sizeof T);
      orc_project2d(K, T, X, uv);
      if (uv[0] < x_min) x_min = uv[0];
      if (uv[0] > x_max) x_max = uv[0];
      if (uv[1] < y_min) y_min = uv[1];
      if (uv[1] > y_max) y_max = uv[1];
    }
  }
  out[0] = x_min; out[1] = y_min; out[2] = x_max; out[3] = y_max;
}
