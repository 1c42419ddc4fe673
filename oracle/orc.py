"""ctypes binding of the CPU ORACLE (oracle/libpf_oracle.so).

TEST INFRASTRUCTURE ONLY.  Imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference leg.  Never imported by the product package.
Parity status: RNG pinned against libstdc++ (check_rng); everything else "parity unpinned"
(see pf_oracle.h).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpf_oracle.so")
MAX_LED = 8


def build(force=False):
    """Compile the oracle (gcc) if missing or stale."""
    src = [os.path.join(_HERE, f) for f in ("pf_oracle.c", "pf_oracle.h", "p3p_oracle.cpp", "Makefile")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in src)
    if stale or force:
        subprocess.check_call(["make", "-C", _HERE, "libpf_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


class Params(C.Structure):
    _fields_ = [("min_angular_noise", C.c_double), ("max_angular_noise", C.c_double),
                ("min_transition_noise", C.c_double), ("max_transition_noise", C.c_double),
                ("back_projection_pixel_tolerance", C.c_double),
                ("back_projection_pixel_tolerance_PF", C.c_double),
                ("marker_downgrade", C.c_int * 5)]


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_up = C.POINTER(C.c_uint)


class PfIn(C.Structure):
    _fields_ = [("N", C.c_int), ("L", C.c_int), ("D", C.c_int),
                ("obj", _dp), ("img", _dp), ("K", _dp), ("E", _dp),
                ("T_cur", _dp), ("T_pred", _dp), ("P", _dp), ("C", _dp),
                ("isi", C.c_uint), ("dt", C.c_double), ("seed", C.c_uint32),
                ("resampler", C.c_int), ("max_iter", C.c_int), ("dist", _dp)]


class PfOut(C.Structure):
    _fields_ = [("T", _dp), ("w_raw", _dp), ("w", _dp), ("pairs", _up), ("npairs", _ip),
                ("E_new", _dp), ("idx", _ip), ("count", _up),
                ("iterations", C.c_int), ("best_iteration", C.c_int),
                ("highest_prob", C.c_double), ("weight_sum", C.c_double),
                ("most_likely_particle_idx", C.c_int), ("most_likely_pose_idx", C.c_int),
                ("accepted", C.c_int), ("resampled", C.c_int), ("flag_fail", C.c_float),
                ("winner_pose", C.c_double * 16), ("winner_pairs", C.c_uint * (2 * MAX_LED)),
                ("winner_npairs", C.c_int), ("isi_out", C.c_uint), ("rng_steps", C.c_uint64)]


class Tracker(C.Structure):
    _fields_ = [("N", C.c_int), ("L", C.c_int), ("obj", C.c_double * (4 * MAX_LED)),
                ("K", C.c_double * 9), ("prm", Params),
                ("current_pose", C.c_double * 16), ("previous_pose", C.c_double * 16),
                ("predicted_pose", C.c_double * 16),
                ("current_time", C.c_double), ("previous_time", C.c_double), ("predicted_time", C.c_double),
                ("pose_covariance", C.c_double * 36), ("isi", C.c_uint),
                ("E", _dp), ("T", _dp), ("w", _dp), ("resampler", C.c_int), ("pf", PfOut)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_rng_skip.restype = C.c_uint32
        L.orc_rng_skip.argtypes = [C.c_uint32, C.c_uint64]
        L.orc_rng_canonical.restype = C.c_double
        L.orc_rng_uniform.restype = C.c_double
        L.orc_rng_uniform.argtypes = [C.c_void_p, C.c_double, C.c_double]
        L.orc_rng_next.restype = C.c_uint32
        L.orc_score.restype = C.c_double
        L.orc_score.argtypes = [C.POINTER(Params), C.c_int, C.c_int, _dp, _dp, _up, _ip]
        L.orc_pf_update.restype = C.c_int
        L.orc_pf_update.argtypes = [C.POINTER(Params), C.POINTER(PfIn), C.POINTER(PfOut)]
        L.orc_resample.argtypes = [C.c_int, _dp, _dp, C.c_int, _ip, _up]
        L.orc_optimise_pose.restype = C.c_int
        L.orc_optimise_pose.argtypes = [_dp, C.c_int, _dp, C.c_int, _dp, _up, C.c_int, _dp, _dp]
        L.orc_predict_pose.argtypes = [_dp, _dp, C.c_double, C.c_double, C.c_double, _dp, _dp]
        L.orc_exponential_map.argtypes = [_dp, _dp]
        L.orc_logarithm_map.argtypes = [_dp, _dp]
        L.orc_project2d.argtypes = [_dp, _dp, _dp, _dp]
        L.orc_tracker_init.restype = C.c_int
        L.orc_tracker_init.argtypes = [C.POINTER(Tracker), C.c_int, C.c_int, _dp, _dp, C.POINTER(Params),
                                       _dp, C.c_double, _dp, C.c_double, C.c_int]
        L.orc_tracker_free.argtypes = [C.POINTER(Tracker)]
        L.orc_tracker_frame.restype = C.c_float
        L.orc_tracker_frame.argtypes = [C.POINTER(Tracker), C.c_double, C.c_int, _dp, C.c_uint32]
        L.orc_weighted_mean_cov.argtypes = [C.c_int, _dp, _dp, _dp, _dp]
        L.orc_distort_point.argtypes = [_dp, _dp, _dp, _dp]
        L.orc_predict_roi.argtypes = [C.c_int, _dp, _dp, _dp, _dp, _dp, C.c_int, _dp, _dp]
        L.orc_mat4_inv.argtypes = [_dp, _dp]
        L.orc_p3p_solve_quartic.argtypes = [_dp, _dp]
        L.orc_p3p_compute_poses.restype = C.c_int
        L.orc_p3p_compute_poses.argtypes = [_dp, _dp, _dp]
        L.orc_image_vectors.argtypes = [_dp, C.c_int, _dp, _dp]
        L.orc_p3p_histogram.restype = C.c_int
        L.orc_p3p_histogram.argtypes = [_dp, C.c_int, _dp, C.c_int, _dp, C.c_double, _up, C.POINTER(C.c_uint64)]
        L.orc_combinations.restype = C.c_int
        L.orc_combinations.argtypes = [C.c_uint, C.c_uint, _up, C.c_int]
        L.orc_check_correspondences.restype = C.c_int
        L.orc_check_correspondences.argtypes = [_dp, C.c_int, _dp, C.c_int, _dp, C.c_int, _up, C.c_double, C.c_double, _ip, _ip,
                                                _ip, _dp, _dp]
        L.orc_p3p_histogram_mode.restype = C.c_int
        L.orc_p3p_histogram_mode.argtypes = [_dp, C.c_int, _dp, C.c_int, _dp, C.c_double, C.c_int, _up,
                                             C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _c64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def make_params(p):
    """p: dict with the reference's parameter names (see scenes.DEFAULT_PARAMS)."""
    q = Params()
    q.min_angular_noise = p["minAngularNoise"]
    q.max_angular_noise = p["maxAngularNoise"]
    q.min_transition_noise = p["minTransitionNoise"]
    q.max_transition_noise = p["maxTransitionNoise"]
    q.back_projection_pixel_tolerance = p["back_projection_pixel_tolerance"]
    q.back_projection_pixel_tolerance_PF = p["back_projection_pixel_tolerance_PF"]
    for i in range(5):
        q.marker_downgrade[i] = int(p.get("bMarkerDowngrade", [0] * 5)[i])
    return q


def obj_h(markers):
    m = np.asarray(markers, dtype=np.float64).reshape(-1, 3)
    return np.ascontiguousarray(np.concatenate([m, np.ones((m.shape[0], 1))], axis=1))


def colmajor(T):
    """(..., 4, 4) row-indexed numpy matrices -> (..., 16) column-major (Eigen layout)."""
    T = np.asarray(T, dtype=np.float64)
    return np.ascontiguousarray(np.swapaxes(T, -1, -2).reshape(T.shape[:-2] + (16,)))


def from_colmajor(a):
    a = np.asarray(a, dtype=np.float64)
    return np.swapaxes(a.reshape(a.shape[:-1] + (4, 4)), -1, -2).copy()


def score(params, img, proj):
    L = proj.shape[0]
    D = img.shape[0]
    img = _c64(img)
    proj = _c64(proj)
    pairs = np.zeros(2 * MAX_LED, dtype=np.uint32)
    n = C.c_int(0)
    prm = make_params(params)
    v = lib().orc_score(C.byref(prm), L, D, _d(img), _d(proj), pairs.ctypes.data_as(_up), C.byref(n))
    return v, pairs[: 2 * n.value].reshape(-1, 2).copy()


def pf_update(params, markers, K, img, E, T_cur, T_pred, P, Cm, isi, dt, seed, resampler=1, max_iter=80, dist=None):
    """One PF update.  All poses are (4,4) numpy matrices (E: (N,4,4)).  Returns a dict."""
    obj = obj_h(markers)
    L = obj.shape[0]
    img = _c64(img).reshape(-1, 2)
    D = img.shape[0]
    Kc = _c64(K).reshape(9)
    Ec = colmajor(E)
    N = Ec.shape[0]
    tc, tp, pc, cc = colmajor(T_cur), colmajor(T_pred), colmajor(P), colmajor(Cm)
    prm = make_params(params)
    i = PfIn()
    i.N, i.L, i.D = N, L, D
    i.obj, i.img, i.K, i.E = _d(obj), _d(img), _d(Kc), _d(Ec)
    i.T_cur, i.T_pred, i.P, i.C = _d(tc), _d(tp), _d(pc), _d(cc)
    i.isi, i.dt, i.seed, i.resampler, i.max_iter = isi, dt, seed & 0xFFFFFFFF, resampler, max_iter
    dc = None if dist is None else _c64(dist).reshape(5)
    i.dist = None if dc is None else _d(dc)
    o = PfOut()
    T = np.zeros((N, 16)); w_raw = np.zeros(N); w = np.zeros(N)
    pairs = np.zeros((N, 2 * MAX_LED), dtype=np.uint32); npairs = np.zeros(N, dtype=np.int32)
    E_new = np.zeros((N, 16)); idx = np.zeros(N, dtype=np.int32); count = np.zeros(N, dtype=np.uint32)
    o.T, o.w_raw, o.w = _d(T), _d(w_raw), _d(w)
    o.pairs, o.npairs = pairs.ctypes.data_as(_up), npairs.ctypes.data_as(_ip)
    o.E_new, o.idx, o.count = _d(E_new), idx.ctypes.data_as(_ip), count.ctypes.data_as(_up)
    rc = lib().orc_pf_update(C.byref(prm), C.byref(i), C.byref(o))
    assert rc == 0
    return dict(T=from_colmajor(T), w_raw=w_raw, w=w, pairs=pairs, npairs=npairs,
                E_new=from_colmajor(E_new), idx=idx, count=count,
                iterations=o.iterations, best_iteration=o.best_iteration,
                highest_prob=o.highest_prob, weight_sum=o.weight_sum,
                most_likely_particle_idx=o.most_likely_particle_idx,
                most_likely_pose_idx=o.most_likely_pose_idx, accepted=o.accepted,
                resampled=o.resampled, flag_fail=o.flag_fail,
                winner_pose=from_colmajor(np.array(o.winner_pose[:])),
                winner_pairs=np.array(o.winner_pairs[: 2 * o.winner_npairs], dtype=np.uint32).reshape(-1, 2),
                isi_out=o.isi_out, rng_steps=o.rng_steps)


def resample(w, r, literal=False):
    w = _c64(w); r = _c64(r)
    N = w.shape[0]
    idx = np.zeros(N, dtype=np.int32); count = np.zeros(N, dtype=np.uint32)
    lib().orc_resample(N, _d(w), _d(r), int(literal), idx.ctypes.data_as(_ip), count.ctypes.data_as(_up))
    return idx, count


def resample_uniforms(seed, N, rng_steps_before):
    """The N uniforms r_k the reference draws at :671, given the stream position."""
    g = C.c_uint32(lib().orc_rng_skip(seed_state(seed), rng_steps_before))
    out = np.empty(N)
    for k in range(N):
        out[k] = lib().orc_rng_uniform(C.byref(g), 0.0, 1.0)
    return out


def seed_state(seed):
    x = (seed & 0xFFFFFFFF) % 2147483647
    return x if x else 1


def optimise_pose(K, markers, img, pairs, pose):
    obj = obj_h(markers)
    img = _c64(img).reshape(-1, 2)
    pr = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
    p = colmajor(pose).copy()
    cov = np.zeros(36)
    Kc = _c64(K).reshape(9)
    it = lib().orc_optimise_pose(_d(Kc), obj.shape[0], _d(obj), img.shape[0], _d(img),
                                 pr.ctypes.data_as(_up), pr.shape[0], _d(p), _d(cov))
    return from_colmajor(p), cov.reshape(6, 6).T.copy(), it


def predict_pose(prev, cur, t_prev, t_cur, t_pred):
    P = np.zeros(16); pred = np.zeros(16)
    a, b = colmajor(prev), colmajor(cur)
    lib().orc_predict_pose(_d(a), _d(b), t_prev, t_cur, t_pred, _d(P), _d(pred))
    return from_colmajor(P), from_colmajor(pred)


def exponential_map(tw):
    tw = _c64(tw); T = np.zeros(16)
    lib().orc_exponential_map(_d(tw), _d(T))
    return from_colmajor(T)


def logarithm_map(T):
    t = colmajor(T); xi = np.zeros(6)
    lib().orc_logarithm_map(_d(t), _d(xi))
    return xi


def project2d(K, T, X):
    Kc = _c64(K).reshape(9); t = colmajor(T); x = _c64(X); uv = np.zeros(2)
    lib().orc_project2d(_d(Kc), _d(t), _d(x), _d(uv))
    return uv


def distort_points(K, D5, pts):
    """LEDDetector::distortPoints on an (n, 2) array."""
    pts = _c64(pts).reshape(-1, 2); out = np.zeros_like(pts); Kc = _c64(K).reshape(9); d = _c64(D5).reshape(5)
    for a, b in zip(pts, out):
        lib().orc_distort_point(_d(Kc), _d(d), _d(a), _d(b))
    return out


def predict_roi(E, Cm, P, pred, markers, K):
    """(x_min, y_min, x_max, y_max) of predictMarkerPositionsInImage + determineROI's extreme-value loop."""
    e = colmajor(E); m = _c64(markers)[:, :3].copy(); out = np.zeros(4)
    lib().orc_predict_roi(e.shape[0], _d(e), _d(colmajor(Cm)), _d(colmajor(P)), _d(colmajor(pred)), _d(m), m.shape[0],
                          _d(_c64(K).reshape(9)), _d(out))
    return out


def p3p_solve_quartic(factors):
    """P3P::solveQuartic (p3p.cpp:238-292): real parts of the four roots."""
    f = _c64(factors).reshape(5); r = np.zeros(4)
    lib().orc_p3p_solve_quartic(_d(f), _d(r))
    return r


def p3p_compute_poses(feature_vectors, world_points):
    """P3P::computePoses (p3p.cpp:65-236).  feature_vectors / world_points: (3, 3) arrays whose COLUMNS are the
    three unit bearing vectors / world points.  Returns (rc, solutions (4, 3, 4) = [R | C])."""
    fv = np.ascontiguousarray(np.asarray(feature_vectors, dtype=np.float64).T).reshape(9)   # column-major
    wp = np.ascontiguousarray(np.asarray(world_points, dtype=np.float64).T).reshape(9)
    sol = np.zeros(48)
    rc = lib().orc_p3p_compute_poses(_d(fv), _d(wp), _d(sol))
    return rc, sol.reshape(4, 4, 3).transpose(0, 2, 1).copy()


def image_vectors(K, img):
    """calculateImageVectors (pose_estimator.cpp:1072-1085): (D, 3) unit bearing vectors."""
    im = _c64(img).reshape(-1, 2); out = np.zeros((im.shape[0], 3))
    lib().orc_image_vectors(_d(_c64(K).reshape(9)), im.shape[0], _d(im), _d(out))
    return out


def p3p_histogram(K, markers, img, tol, dup_mode=0):
    """hist_corr of PoseEstimator::initialise (pose_estimator.cpp:1524-1717): (D, L) uint32 votes and the counters
    (triples, p3p_solved, solutions_projected, solutions_voted).  dup_mode: 0 the reference's duplicate-solution
    rule (:1659-1663), 1 never drop (upper bracket), 2 drop every near-duplicate (lower bracket)."""
    im = _c64(img).reshape(-1, 2); ob = obj_h(np.asarray(markers)[:, :3])
    D, L = im.shape[0], ob.shape[0]
    hist = np.zeros((D, L), dtype=np.uint32); st = (C.c_uint64 * 4)()
    lib().orc_p3p_histogram_mode(_d(_c64(K).reshape(9)), L, _d(ob), D, _d(im), float(tol), int(dup_mode),
                                 hist.ctypes.data_as(_up), st)
    return hist, tuple(int(x) for x in st)


def combinations(n, k):
    """Combinations::combinationsNoReplacement(n, k) (combinations.cpp:52-128): rows of 1-based indices."""
    from math import comb
    out = np.zeros((comb(n, k), k), dtype=np.uint32)
    rows = lib().orc_combinations(n, k, out.ctypes.data_as(_up), out.shape[0])
    return out[:rows]


def check_correspondences(K, markers, img, corr, tol, certainty_threshold):
    """Per-combination body of checkCorrespondences (pose_estimator.cpp:1340-1445) for one correspondence set
    ((n, 2) 1-based (LED, detection) pairs): list of dicts (rc, valid, best, squared_error, pose)."""
    im = _c64(img).reshape(-1, 2); ob = obj_h(np.asarray(markers)[:, :3])
    c = np.ascontiguousarray(np.asarray(corr, dtype=np.uint32).reshape(-1, 2))
    cap = 1024
    rc = np.zeros(cap, dtype=np.int32); valid = np.zeros(cap, dtype=np.int32); best = np.zeros(cap, dtype=np.int32)
    err = np.zeros(cap); pose = np.zeros((cap, 16))
    nc = lib().orc_check_correspondences(_d(_c64(K).reshape(9)), ob.shape[0], _d(ob), im.shape[0], _d(im), c.shape[0],
                                         c.ctypes.data_as(_up), float(tol), float(certainty_threshold),
                                         rc.ctypes.data_as(_ip), valid.ctypes.data_as(_ip), best.ctypes.data_as(_ip), _d(err), _d(pose))
    return [dict(rc=int(rc[i]), valid=int(valid[i]), best=int(best[i]), squared_error=float(err[i]), pose=from_colmajor(pose[i]))
            for i in range(nc)]


def weighted_mean_cov(T, w):
    t = colmajor(T); w = _c64(w)
    mp = np.zeros(16); cov = np.zeros(36)
    lib().orc_weighted_mean_cov(t.shape[0], _d(t), _d(w), _d(mp), _d(cov))
    return from_colmajor(mp), cov.reshape(6, 6).T.copy()


class OracleTracker:
    """estimateBodyPose PF branch without detection/ROI (orc_tracker_*)."""

    def __init__(self, N, markers, K, params, pose0, t0, pose1, t1, resampler=1):
        self.t = Tracker()
        obj = obj_h(markers)
        Kc = _c64(K).reshape(9)
        prm = make_params(params)
        p0, p1 = colmajor(pose0), colmajor(pose1)
        rc = lib().orc_tracker_init(C.byref(self.t), N, obj.shape[0], _d(obj), _d(Kc), C.byref(prm),
                                    _d(p0), t0, _d(p1), t1, resampler)
        assert rc == 0
        self.N = N

    def frame(self, time_to_predict, img, seed):
        img = _c64(img).reshape(-1, 2)
        return lib().orc_tracker_frame(C.byref(self.t), time_to_predict, img.shape[0], _d(img), seed & 0xFFFFFFFF)

    @property
    def current_pose(self):
        return from_colmajor(np.array(self.t.current_pose[:]))

    @property
    def predicted_pose(self):
        return from_colmajor(np.array(self.t.predicted_pose[:]))

    @property
    def covariance(self):
        return np.array(self.t.pose_covariance[:]).reshape(6, 6).T.copy()

    @property
    def pf(self):
        return self.t.pf

    def particles(self):
        return from_colmajor(np.ctypeslib.as_array(self.t.T, shape=(self.N, 16)).copy())

    def resampled(self):
        return from_colmajor(np.ctypeslib.as_array(self.t.E, shape=(self.N, 16)).copy())

    def weights_raw(self):
        return np.ctypeslib.as_array(self.t.pf.w_raw, shape=(self.N,)).copy()

    def close(self):
        if self.t.E:
            lib().orc_tracker_free(C.byref(self.t))

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
