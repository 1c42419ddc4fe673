"""ctypes binding of libpfmpe.so (the C ABI in include/pfmpe.h).

Thin, mechanical mirror of the header: tests and bench.py call the CUDA path through these
functions only.  Loading fails loudly when the library has not been built; creating a handle
fails loudly (PfmpeError) when no CUDA device is usable -- there is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PFMPE_LIB", os.path.join(_HERE, "libpfmpe.so"))   # override: A/B experiments only
MAX_LED = 8

OK, ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_STATE, ERR_UNSUPPORTED, ERR_COMM = 0, -1, -2, -3, -4, -5, -6
FLAG_MEAN_COV = 1
FLAG_DISTORT = 2

EXPORTS = [
    "pfmpe_create", "pfmpe_destroy", "pfmpe_strerror", "pfmpe_last_error", "pfmpe_version",
    "pfmpe_set_markers", "pfmpe_set_camera", "pfmpe_set_params", "pfmpe_set_particles",
    "pfmpe_update", "pfmpe_update_batch",
    "pfmpe_get_particles", "pfmpe_get_weights", "pfmpe_get_resample_indices", "pfmpe_get_resample_counts",
    "pfmpe_optimise_pose", "pfmpe_predict_pose", "pfmpe_exponential_map", "pfmpe_logarithm_map",
    "pfmpe_track_init", "pfmpe_track_set_state", "pfmpe_track_upload", "pfmpe_track_attach", "pfmpe_track_run", "pfmpe_track_replay", "pfmpe_track_results",
    "pfmpe_track_frame_host", "pfmpe_track_frame_host_batch", "pfmpe_track_sequence_host",
    "pfmpe_track_frame_device_batch", "pfmpe_track_sequence_device",
    "pfmpe_predict_roi", "pfmpe_p3p_histogram", "pfmpe_p3p_histogram_dev", "pfmpe_p3p_histogram_static", "pfmpe_p3p_solve", "pfmpe_p3p_check",
    "pfmpe_shard_configure", "pfmpe_shard_begin_frame", "pfmpe_shard_score", "pfmpe_shard_resample",
    "pfmpe_shard_export", "pfmpe_shard_keep", "pfmpe_shard_import", "pfmpe_shard_winner",
    "pfmpe_shard_mail_create", "pfmpe_shard_mail_set_peers", "pfmpe_shard_collect", "pfmpe_shard_advance",
    "pfmpe_shard_barrier", "pfmpe_shard_signal", "pfmpe_shard_wait", "pfmpe_shard_winner_publish",
    "pfmpe_shard_frame", "pfmpe_shard_track_frame_host", "pfmpe_shard_passes", "pfmpe_state_checksum",
    "pfmpe_shard_local_ptr", "pfmpe_shard_ipc_export", "pfmpe_shard_ipc_open", "pfmpe_shard_set_peers", "pfmpe_shard_scatter",
    "pfmpe_set_stream", "pfmpe_sync", "pfmpe_last_launch_ms", "pfmpe_launch_count", "pfmpe_launch_info", "pfmpe_debug_timeline",
]


class PfmpeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("pfmpe error %d: %s" % (code, msg))
        self.code = code


class Config(C.Structure):
    _fields_ = [("n_particles", C.c_int32), ("n_trackers", C.c_int32), ("n_leds", C.c_int32),
                ("max_blobs", C.c_int32), ("device", C.c_int32), ("flags", C.c_uint32),
                ("reserved", C.c_int32 * 4)]


class PfParams(C.Structure):
    _fields_ = [("min_angular_noise", C.c_double), ("max_angular_noise", C.c_double),
                ("min_transition_noise", C.c_double), ("max_transition_noise", C.c_double),
                ("back_projection_pixel_tolerance", C.c_double),
                ("back_projection_pixel_tolerance_PF", C.c_double),
                ("marker_downgrade", C.c_int32 * 5), ("reserved", C.c_int32)]


class FrameIn(C.Structure):
    _fields_ = [("image_points", C.POINTER(C.c_double)), ("n_image_points", C.c_int32),
                ("it_since_initialized", C.c_uint32),
                ("current_pose", C.c_double * 16), ("predicted_pose", C.c_double * 16),
                ("prediction_matrix", C.c_double * 16), ("cam_move_inv", C.c_double * 16),
                ("dt", C.c_double), ("seed", C.c_uint32), ("reserved", C.c_uint32)]


class FrameOut(C.Structure):
    _fields_ = [("status", C.c_int32), ("iterations", C.c_int32), ("best_iteration", C.c_int32),
                ("resampled", C.c_int32), ("highest_prob", C.c_double), ("weight_sum", C.c_double),
                ("most_likely_particle_idx", C.c_int32), ("most_likely_pose_idx", C.c_int32),
                ("max_count", C.c_uint32), ("n_correspondences", C.c_int32),
                ("correspondences", C.c_uint32 * (2 * MAX_LED)),
                ("winner_pose", C.c_double * 16), ("mean_pose", C.c_double * 16), ("mean_cov", C.c_double * 36)]


class P3pStats(C.Structure):
    _fields_ = [("triples", C.c_uint64), ("p3p_solved", C.c_uint64), ("solutions_projected", C.c_uint64),
                ("solutions_voted", C.c_uint64), ("device_ms", C.c_float), ("reserved", C.c_int32)]


class P3pCheckResult(C.Structure):
    _fields_ = [("rc", C.c_int32), ("valid", C.c_int32), ("best", C.c_int32), ("reserved", C.c_int32),
                ("squared_error", C.c_double), ("pose", C.c_double * 16)]


P3P_CHECK_STRIDE = 56


class TrackResult(C.Structure):
    _fields_ = [("status", C.c_int32), ("flag_fail", C.c_float), ("iterations", C.c_int32),
                ("gn_iterations", C.c_int32), ("most_likely_pose_idx", C.c_int32), ("max_count", C.c_uint32),
                ("highest_prob", C.c_double), ("weight_sum", C.c_double),
                ("pose", C.c_double * 16), ("cov", C.c_double * 36), ("winner_pose", C.c_double * 16),
                ("n_correspondences", C.c_int32), ("correspondences", C.c_uint32 * (2 * MAX_LED)),
                ("reserved", C.c_int32)]


_lib = None
_dp = C.POINTER(C.c_double)


def lib():
    """Load libpfmpe.so; raises if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("libpfmpe.so is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(make -C pf_monocular_pose_estimator_b200/csrc). There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        L.pfmpe_strerror.restype = C.c_char_p
        L.pfmpe_last_error.restype = C.c_char_p
        L.pfmpe_last_error.argtypes = [C.c_void_p]
        L.pfmpe_create.argtypes = [C.POINTER(Config), C.POINTER(C.c_void_p)]
        L.pfmpe_destroy.argtypes = [C.c_void_p]
        L.pfmpe_set_markers.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int]
        L.pfmpe_set_camera.argtypes = [C.c_void_p, _dp, _dp]
        L.pfmpe_set_params.argtypes = [C.c_void_p, C.POINTER(PfParams)]
        L.pfmpe_set_particles.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int]
        L.pfmpe_update.argtypes = [C.c_void_p, C.c_int, C.POINTER(FrameIn), C.POINTER(FrameOut)]
        L.pfmpe_update_batch.argtypes = [C.c_void_p, C.POINTER(FrameIn), C.POINTER(FrameOut)]
        L.pfmpe_get_particles.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
        L.pfmpe_get_weights.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp]
        L.pfmpe_get_resample_indices.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int32)]
        L.pfmpe_get_resample_counts.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint32)]
        L.pfmpe_optimise_pose.argtypes = [_dp, _dp, C.c_int, _dp, C.c_int, C.POINTER(C.c_uint32), C.c_int, _dp, _dp,
                                          C.POINTER(C.c_int)]
        L.pfmpe_predict_pose.argtypes = [_dp, _dp, C.c_double, C.c_double, C.c_double, _dp, _dp]
        L.pfmpe_exponential_map.argtypes = [_dp, _dp]
        L.pfmpe_logarithm_map.argtypes = [_dp, _dp]
        L.pfmpe_track_init.argtypes = [C.c_void_p, C.c_int, _dp, C.c_double, _dp, C.c_double]
        L.pfmpe_track_set_state.argtypes = [C.c_void_p, C.c_int, _dp, C.c_double, _dp, C.c_double, C.c_uint32]
        L.pfmpe_track_upload.argtypes = [C.c_void_p, C.c_int, _dp, C.POINTER(C.c_int32), _dp, C.POINTER(C.c_uint32)]
        L.pfmpe_track_attach.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.pfmpe_track_run.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.pfmpe_track_replay.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.pfmpe_track_results.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(TrackResult)]
        L.pfmpe_track_frame_host.argtypes = [C.c_void_p, C.c_int, C.c_double, _dp, C.c_int32, C.c_uint32, C.POINTER(TrackResult)]
        L.pfmpe_track_frame_host_batch.argtypes = [C.c_void_p, C.c_double, _dp, C.POINTER(C.c_int32), C.POINTER(C.c_uint32),
                                                   C.POINTER(TrackResult)]
        L.pfmpe_track_sequence_host.argtypes = [C.c_void_p, C.c_int, _dp, _dp, C.POINTER(C.c_int32), C.POINTER(C.c_uint32),
                                                C.POINTER(TrackResult), _dp]
        L.pfmpe_track_sequence_device.argtypes = L.pfmpe_track_sequence_host.argtypes
        L.pfmpe_track_frame_device_batch.argtypes = L.pfmpe_track_frame_host_batch.argtypes
        L.pfmpe_predict_roi.argtypes = [C.c_void_p, C.c_int, _dp, _dp, _dp, _dp]
        L.pfmpe_p3p_histogram.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int32, C.POINTER(C.c_uint32), C.POINTER(P3pStats)]
        L.pfmpe_p3p_histogram_dev.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int32, C.POINTER(C.c_uint32),
                                              C.POINTER(P3pStats)]
        L.pfmpe_p3p_histogram_static.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int32, C.POINTER(C.c_uint32), C.POINTER(P3pStats)]
        L.pfmpe_p3p_solve.argtypes = [C.c_void_p, C.c_int32, _dp, _dp, _dp, C.POINTER(C.c_int32)]
        L.pfmpe_p3p_check.argtypes = [C.c_void_p, C.c_int, _dp, C.c_int32, C.POINTER(C.c_uint32), C.POINTER(C.c_int32), C.c_int32,
                                      C.c_int32, C.c_double, C.POINTER(P3pCheckResult)]
        _i3 = C.POINTER(C.c_int64)
        L.pfmpe_shard_configure.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
        L.pfmpe_shard_begin_frame.argtypes = [C.c_void_p, C.POINTER(FrameIn)]
        L.pfmpe_shard_score.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.pfmpe_shard_resample.argtypes = [C.c_void_p, C.c_int64, C.c_int, _i3, _i3, C.c_void_p]
        L.pfmpe_shard_export.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
        L.pfmpe_shard_import.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        L.pfmpe_shard_keep.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64]
        L.pfmpe_shard_local_ptr.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.pfmpe_shard_ipc_export.argtypes = [C.c_void_p, C.c_char_p]
        L.pfmpe_shard_ipc_open.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]
        L.pfmpe_shard_set_peers.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        L.pfmpe_shard_scatter.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
        L.pfmpe_shard_mail_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_void_p), C.c_char_p]
        L.pfmpe_shard_mail_set_peers.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
        L.pfmpe_shard_collect.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
        L.pfmpe_shard_advance.argtypes = [C.c_void_p]
        L.pfmpe_shard_barrier.argtypes = [C.c_void_p]
        L.pfmpe_shard_signal.argtypes = [C.c_void_p]
        L.pfmpe_shard_wait.argtypes = [C.c_void_p]
        L.pfmpe_shard_frame.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(FrameIn), C.POINTER(FrameOut)]
        L.pfmpe_shard_track_frame_host.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_double, _dp, C.c_int32, C.c_uint32,
                                                   C.POINTER(TrackResult)]
        L.pfmpe_shard_passes.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.pfmpe_state_checksum.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
        L.pfmpe_shard_winner_publish.argtypes = [C.c_void_p, C.c_int]
        L.pfmpe_shard_winner.argtypes = [C.c_void_p, C.c_int, C.POINTER(FrameOut)]
        L.pfmpe_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.pfmpe_sync.argtypes = [C.c_void_p]
        L.pfmpe_last_launch_ms.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
        L.pfmpe_launch_count.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.pfmpe_launch_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
        L.pfmpe_debug_timeline.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _c64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def colmajor(T):
    """(..., 4, 4) numpy matrices -> (..., 16) column-major doubles (Eigen::Matrix4d layout)."""
    T = np.asarray(T, dtype=np.float64)
    return np.ascontiguousarray(np.swapaxes(T, -1, -2).reshape(T.shape[:-2] + (16,)))


def from_colmajor(a):
    a = np.asarray(a, dtype=np.float64)
    return np.swapaxes(a.reshape(a.shape[:-1] + (4, 4)), -1, -2).copy()


def make_params(p):
    q = PfParams()
    q.min_angular_noise = p["minAngularNoise"]
    q.max_angular_noise = p["maxAngularNoise"]
    q.min_transition_noise = p["minTransitionNoise"]
    q.max_transition_noise = p["maxTransitionNoise"]
    q.back_projection_pixel_tolerance = p["back_projection_pixel_tolerance"]
    q.back_projection_pixel_tolerance_PF = p["back_projection_pixel_tolerance_PF"]
    dg = p.get("bMarkerDowngrade", [0] * 5)
    for i in range(5):
        q.marker_downgrade[i] = int(dg[i])
    return q


def make_frame_in(img, T_cur, T_pred, P, Cm, isi, dt, seed, keep):
    """Build a pfmpe_frame_in; `keep` collects the numpy buffers that must outlive the call."""
    img = _c64(img).reshape(-1, 2)
    keep.append(img)
    fi = FrameIn()
    fi.image_points = _d(img)
    fi.n_image_points = img.shape[0]
    fi.it_since_initialized = isi
    for name, M in (("current_pose", T_cur), ("predicted_pose", T_pred), ("prediction_matrix", P), ("cam_move_inv", Cm)):
        arr = colmajor(M)
        for i in range(16):
            getattr(fi, name)[i] = arr[i]
    fi.dt = dt
    fi.seed = seed & 0xFFFFFFFF
    return fi


def unpack_frame_out(o):
    nc = o.n_correspondences
    return dict(status=o.status, iterations=o.iterations, best_iteration=o.best_iteration, resampled=o.resampled,
                highest_prob=o.highest_prob, weight_sum=o.weight_sum,
                most_likely_particle_idx=o.most_likely_particle_idx, most_likely_pose_idx=o.most_likely_pose_idx,
                max_count=o.max_count,
                correspondences=np.array(o.correspondences[: 2 * nc], dtype=np.uint32).reshape(-1, 2),
                winner_pose=from_colmajor(np.array(o.winner_pose[:])),
                mean_pose=from_colmajor(np.array(o.mean_pose[:])),
                mean_cov=np.array(o.mean_cov[:]).reshape(6, 6).T.copy())


class Handle:
    """RAII wrapper of a pfmpe_handle."""

    def __init__(self, n_particles, n_trackers=1, n_leds=5, max_blobs=32, device=0, flags=0, force_mode=0, grid_cap=0):
        self._h = C.c_void_p()
        cfg = Config()
        cfg.n_particles, cfg.n_trackers, cfg.n_leds, cfg.max_blobs = n_particles, n_trackers, n_leds, max_blobs
        cfg.device, cfg.flags = device, flags
        cfg.reserved[0], cfg.reserved[1] = force_mode, grid_cap
        rc = lib().pfmpe_create(C.byref(cfg), C.byref(self._h))
        if rc != OK:
            msg = lib().pfmpe_last_error(None).decode() or lib().pfmpe_strerror(rc).decode()
            self._h = C.c_void_p()
            raise PfmpeError(rc, msg)
        self.N, self.T, self.L, self.max_blobs = n_particles, n_trackers, n_leds, max_blobs

    def _chk(self, rc):
        if rc != OK:
            raise PfmpeError(rc, (lib().pfmpe_last_error(self._h) or b"").decode() + " (" + lib().pfmpe_strerror(rc).decode() + ")")

    def close(self):
        if self._h:
            lib().pfmpe_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # configuration
    def set_markers(self, xyz, tracker=-1):
        a = _c64(xyz).reshape(-1, 3)
        self._chk(lib().pfmpe_set_markers(self._h, tracker, _d(a), a.shape[0]))

    def set_camera(self, K, D=None):
        k = _c64(K).reshape(9)
        d = _c64(D).reshape(-1) if D is not None else None
        self._chk(lib().pfmpe_set_camera(self._h, _d(k), _d(d) if d is not None else None))

    def set_params(self, params):
        q = make_params(params)
        self._chk(lib().pfmpe_set_params(self._h, C.byref(q)))

    def set_particles(self, poses, tracker=0):
        a = colmajor(poses)
        self._chk(lib().pfmpe_set_particles(self._h, tracker, _d(a), a.shape[0]))

    def configure(self, markers, K, params, D5=None):
        self.set_markers(markers)
        self.set_camera(K, D5)
        self.set_params(params)

    # hot path
    def update(self, img, T_cur, T_pred, P, Cm, isi, dt, seed, tracker=0):
        keep = []
        fi = make_frame_in(img, T_cur, T_pred, P, Cm, isi, dt, seed, keep)
        fo = FrameOut()
        self._chk(lib().pfmpe_update(self._h, tracker, C.byref(fi), C.byref(fo)))
        return unpack_frame_out(fo)

    def update_batch(self, frames):
        """frames: list (len n_trackers) of dicts with keys img,T_cur,T_pred,P,Cm,isi,dt,seed."""
        keep = []
        arr_in = (FrameIn * self.T)()
        for t, f in enumerate(frames):
            arr_in[t] = make_frame_in(f["img"], f["T_cur"], f["T_pred"], f["P"], f["Cm"], f["isi"], f["dt"], f["seed"], keep)
        arr_out = (FrameOut * self.T)()
        self._chk(lib().pfmpe_update_batch(self._h, arr_in, arr_out))
        return [unpack_frame_out(arr_out[t]) for t in range(self.T)]

    # getters
    def particles(self, which=0, tracker=0):
        out = np.zeros((self.N, 16))
        self._chk(lib().pfmpe_get_particles(self._h, tracker, which, _d(out)))
        return from_colmajor(out)

    def weights(self, normalised=False, tracker=0):
        out = np.zeros(self.N)
        self._chk(lib().pfmpe_get_weights(self._h, tracker, int(normalised), _d(out)))
        return out

    def resample_indices(self, tracker=0):
        out = np.zeros(self.N, dtype=np.int32)
        self._chk(lib().pfmpe_get_resample_indices(self._h, tracker, out.ctypes.data_as(C.POINTER(C.c_int32))))
        return out

    def resample_counts(self, tracker=0):
        out = np.zeros(self.N, dtype=np.uint32)
        self._chk(lib().pfmpe_get_resample_counts(self._h, tracker, out.ctypes.data_as(C.POINTER(C.c_uint32))))
        return out

    # device-resident tracking
    def track_init(self, pose0, t0, pose1, t1, tracker=0):
        a, b = colmajor(pose0), colmajor(pose1)
        self._chk(lib().pfmpe_track_init(self._h, tracker, _d(a), t0, _d(b), t1))

    def track_set_state(self, prev_pose, t_prev, cur_pose, t_cur, isi, tracker=0):
        """previous_pose_ / current_pose_ / it_since_initialized_ of a filter seeded with set_particles (:166-186)."""
        self._chk(lib().pfmpe_track_set_state(self._h, tracker, _d(colmajor(prev_pose)), float(t_prev), _d(colmajor(cur_pose)),
                                              float(t_cur), int(isi)))

    def track_upload(self, image_points, counts, times, seeds):
        ip = _c64(image_points)
        cn = np.ascontiguousarray(counts, dtype=np.int32)
        tm = _c64(times)
        sd = np.ascontiguousarray(seeds, dtype=np.uint32)
        n_frames = tm.shape[0]
        assert ip.shape == (self.T, n_frames, self.max_blobs, 2), ip.shape
        assert cn.shape == (self.T, n_frames) and sd.shape == (self.T, n_frames)
        self._chk(lib().pfmpe_track_upload(self._h, n_frames, _d(ip), cn.ctypes.data_as(C.POINTER(C.c_int32)), _d(tm),
                                          sd.ctypes.data_as(C.POINTER(C.c_uint32))))

    def track_attach(self, n_frames, d_image_points, d_counts, d_times, d_seeds):
        self._chk(lib().pfmpe_track_attach(self._h, n_frames, d_image_points, d_counts, d_times, d_seeds))

    def track_run(self, first_frame, n_frames):
        self._chk(lib().pfmpe_track_run(self._h, first_frame, n_frames))

    def track_replay(self, first_frame, n_frames):
        self._chk(lib().pfmpe_track_replay(self._h, first_frame, n_frames))

    def track_frame_host(self, time_to_predict, image_points, seed, tracker=0):
        """One tracked frame through the per-frame C-ABI call with host buffers (host predictPose / Gauss-Newton)."""
        ip = _c64(np.asarray(image_points).reshape(-1, 2))
        r = TrackResult()
        self._chk(lib().pfmpe_track_frame_host(self._h, tracker, float(time_to_predict), _d(ip), ip.shape[0], int(seed),
                                              C.byref(r)))
        return self._unpack_track([r])[0]

    def predict_roi(self, Cm, P, pred, tracker=0):
        """(x_min, y_min, x_max, y_max) over all resampled particles x LEDs + the predicted pose."""
        out = np.zeros(4)
        self._chk(lib().pfmpe_predict_roi(self._h, tracker, _d(colmajor(Cm)), _d(colmajor(P)), _d(colmajor(pred)), _d(out)))
        return out

    def p3p_histogram(self, image_points, tracker=0, d_image_points=None):
        """hist_corr of PoseEstimator::initialise (pose_estimator.cpp:1524-1717): (D, L) uint32 votes + counters.
        d_image_points: device pointer (int) of D x 2 doubles already resident in HBM; image_points then only
        gives the shape."""
        im = _c64(image_points).reshape(-1, 2)
        D, L = im.shape[0], self.L
        hist = np.zeros((D, L), dtype=np.uint32)
        st = P3pStats()
        hp = hist.ctypes.data_as(C.POINTER(C.c_uint32))
        if d_image_points is None:
            self._chk(lib().pfmpe_p3p_histogram(self._h, tracker, _d(im), D, hp, C.byref(st)))
        else:
            self._chk(lib().pfmpe_p3p_histogram_dev(self._h, tracker, C.c_void_p(int(d_image_points)), D, hp, C.byref(st)))
        return hist, dict(triples=int(st.triples), p3p_solved=int(st.p3p_solved),
                          solutions_projected=int(st.solutions_projected), solutions_voted=int(st.solutions_voted),
                          device_ms=float(st.device_ms))

    def p3p_histogram_static(self, image_points, tracker=0):
        """pfmpe_p3p_histogram_static: the same vote with its scratch in the library's file-scope device array (no
        allocation in the call; at most PFMPE_P3P_STATIC_MAX_POINTS = 256 detections)."""
        im = _c64(image_points).reshape(-1, 2)
        D, L = im.shape[0], self.L
        hist = np.zeros((D, L), dtype=np.uint32)
        st = P3pStats()
        self._chk(lib().pfmpe_p3p_histogram_static(self._h, tracker, _d(im), D, hist.ctypes.data_as(C.POINTER(C.c_uint32)), C.byref(st)))
        return hist, dict(triples=int(st.triples), p3p_solved=int(st.p3p_solved),
                          solutions_projected=int(st.solutions_projected), solutions_voted=int(st.solutions_voted),
                          device_ms=float(st.device_ms))

    def p3p_check(self, image_points, correspondence_sets, certainty_threshold, tracker=0):
        """Per-combination body of checkCorrespondences (pose_estimator.cpp:1340-1445) for a list of correspondence
        sets (each an (n, 2) array of 1-based (LED, detection) pairs).  Returns, per set, a list of dicts
        (rc, valid, best, squared_error, pose (4, 4)) in the reference's combination order."""
        im = _c64(image_points).reshape(-1, 2)
        M = len(correspondence_sets)
        corr = np.zeros((M, 16), dtype=np.uint32); cnt = np.zeros(M, dtype=np.int32)
        for i, c in enumerate(correspondence_sets):
            c = np.asarray(c, dtype=np.uint32).reshape(-1, 2)
            cnt[i] = c.shape[0]
            corr[i, :2 * c.shape[0]] = c.reshape(-1)
        nmax = max(3, int(cnt.max())) if M else 3
        stride = nmax * (nmax - 1) * (nmax - 2) // 6
        res = (P3pCheckResult * (M * stride))()
        self._chk(lib().pfmpe_p3p_check(self._h, tracker, _d(im), im.shape[0], corr.ctypes.data_as(C.POINTER(C.c_uint32)),
                                        cnt.ctypes.data_as(C.POINTER(C.c_int32)), M, stride, float(certainty_threshold), res))
        out = []
        for i in range(M):
            n = int(cnt[i]); nc = n * (n - 1) * (n - 2) // 6 if n >= 3 else 0
            out.append([dict(rc=r.rc, valid=r.valid, best=r.best, squared_error=r.squared_error,
                             pose=from_colmajor(np.array(r.pose[:]))) for r in res[i * stride:i * stride + nc]])
        return out

    def p3p_solve(self, feature_vectors, world_points):
        """Batched P3P::computePoses (p3p.cpp:65-236).  feature_vectors, world_points: (n, 3, 3) with the three
        vectors / points as COLUMNS.  Returns rc (n,) and solutions (n, 4, 3, 4) = [R | C]."""
        fv = np.ascontiguousarray(np.asarray(feature_vectors, dtype=np.float64).transpose(0, 2, 1)).reshape(-1, 9)
        wp = np.ascontiguousarray(np.asarray(world_points, dtype=np.float64).transpose(0, 2, 1)).reshape(-1, 9)
        n = fv.shape[0]
        sol = np.zeros((n, 48)); rc = np.zeros(n, dtype=np.int32)
        self._chk(lib().pfmpe_p3p_solve(self._h, n, _d(fv), _d(wp), _d(sol), rc.ctypes.data_as(C.POINTER(C.c_int32))))
        return rc, sol.reshape(n, 4, 4, 3).transpose(0, 1, 3, 2).copy()

    # ---- one filter sharded over several handles / GPUs (device pointers are plain ints) ----
    def shard_configure(self, global_offset, n_global):
        self._chk(lib().pfmpe_shard_configure(self._h, int(global_offset), int(n_global)))

    def shard_begin_frame(self, frame_in):
        self._chk(lib().pfmpe_shard_begin_frame(self._h, C.byref(frame_in)))

    def shard_score(self, iteration, d_partial):
        self._chk(lib().pfmpe_shard_score(self._h, int(iteration), C.c_void_p(int(d_partial))))

    def shard_resample(self, sum_fx, iterations, lower, total, d_partial):
        lo = (C.c_int64 * 3)(*[int(v) for v in lower])
        to = (C.c_int64 * 3)(*[int(v) for v in total])
        self._chk(lib().pfmpe_shard_resample(self._h, int(sum_fx), int(iterations), lo, to, C.c_void_p(int(d_partial))))

    def shard_export(self, first_owned_slot, first_slot, count, d_rows):
        self._chk(lib().pfmpe_shard_export(self._h, int(first_owned_slot), int(first_slot), int(count),
                                           C.c_void_p(int(d_rows))))

    def shard_import(self, d_rows, local_offset, count):
        self._chk(lib().pfmpe_shard_import(self._h, C.c_void_p(int(d_rows)), int(local_offset), int(count)))

    def shard_keep(self, first_owned_slot, first_slot, count, local_offset):
        self._chk(lib().pfmpe_shard_keep(self._h, int(first_owned_slot), int(first_slot), int(count), int(local_offset)))

    def shard_local_ptr(self):
        q = C.c_void_p()
        self._chk(lib().pfmpe_shard_local_ptr(self._h, C.byref(q)))
        return int(q.value)

    def shard_ipc_export(self):
        buf = C.create_string_buffer(64)
        self._chk(lib().pfmpe_shard_ipc_export(self._h, buf))
        return buf.raw

    def shard_ipc_open(self, handle_bytes):
        q = C.c_void_p()
        self._chk(lib().pfmpe_shard_ipc_open(self._h, C.create_string_buffer(bytes(handle_bytes), 64), C.byref(q)))
        return int(q.value)

    def shard_set_peers(self, ptrs):
        arr = (C.c_void_p * max(len(ptrs), 1))(*[int(v) for v in ptrs])
        self._chk(lib().pfmpe_shard_set_peers(self._h, len(ptrs), arr))   # empty list: back to the row exchange

    def shard_scatter(self, first_owned_slot, count):
        self._chk(lib().pfmpe_shard_scatter(self._h, int(first_owned_slot), int(count)))

    def shard_mail_create(self, world, rank):
        """-> (device pointer of this rank's mailbox, its 64-byte CUDA IPC handle)"""
        q = C.c_void_p()
        buf = C.create_string_buffer(64)
        self._chk(lib().pfmpe_shard_mail_create(self._h, world, rank, C.byref(q), buf))
        return int(q.value), buf.raw

    def shard_mail_set_peers(self, ptrs):
        arr = (C.c_void_p * len(ptrs))(*[int(v) for v in ptrs])
        self._chk(lib().pfmpe_shard_mail_set_peers(self._h, arr))

    def shard_collect(self, world, only_rank=-1):
        out = np.zeros((world, 32), dtype=np.int64)
        self._chk(lib().pfmpe_shard_collect(self._h, int(only_rank), out.ctypes.data_as(C.POINTER(C.c_int64))))
        return out

    def shard_advance(self):
        self._chk(lib().pfmpe_shard_advance(self._h))

    def shard_barrier(self):
        self._chk(lib().pfmpe_shard_barrier(self._h))

    def state_checksum(self, tracker=0):
        """Position-keyed device checksums: [PoseParticle, raw weights, newPoseEstimation] (pfmpe_state_checksum)."""
        out = (C.c_uint64 * 3)()
        self._chk(lib().pfmpe_state_checksum(self._h, tracker, out))
        return [int(v) for v in out]

    def shard_passes(self):
        n = C.c_uint64()
        self._chk(lib().pfmpe_shard_passes(self._h, C.byref(n)))
        return int(n.value)

    def shard_winner_publish(self, local_index):
        self._chk(lib().pfmpe_shard_winner_publish(self._h, int(local_index)))

    def shard_winner(self, local_index):
        o = FrameOut()
        self._chk(lib().pfmpe_shard_winner(self._h, int(local_index), C.byref(o)))
        return unpack_frame_out(o)

    def track_sequence_host(self, times, image_points, counts, seeds, device_tail=False, out=None):
        """Per-frame host-buffer tracking of a whole sequence, looped in C (layouts of track_upload).
        device_tail: pfmpe_track_sequence_device (predictPose / Gauss-Newton on the device, row f2) instead of the host tail.
        out: a result array of an earlier call over as many frames, reused (a caller's own, already-touched buffer).
        Returns (results[tracker][frame], seconds per frame)."""
        ip = _c64(image_points)
        cn = np.ascontiguousarray(counts, dtype=np.int32)
        tm = _c64(times)
        sd = np.ascontiguousarray(seeds, dtype=np.uint32)
        F = tm.shape[0]
        assert ip.shape == (self.T, F, self.max_blobs, 2) and cn.shape == (self.T, F) and sd.shape == (self.T, F)
        arr = out if out is not None else (TrackResult * (self.T * F))()
        assert len(arr) == self.T * F
        secs = np.zeros(F)
        fn = lib().pfmpe_track_sequence_device if device_tail else lib().pfmpe_track_sequence_host
        self._chk(fn(self._h, F, _d(tm), _d(ip), cn.ctypes.data_as(C.POINTER(C.c_int32)),
                     sd.ctypes.data_as(C.POINTER(C.c_uint32)), arr, _d(secs)))
        return arr, secs

    def track_frame_device_batch(self, time_to_predict, image_points, counts, seeds):
        """pfmpe_track_frame_device_batch: one tracked frame of every tracker, host buffers in / out, FP64 tail on the device."""
        ip = _c64(image_points)
        cn = np.ascontiguousarray(counts, dtype=np.int32)
        sd = np.ascontiguousarray(seeds, dtype=np.uint32)
        assert ip.shape == (self.T, self.max_blobs, 2) and cn.shape == (self.T,) and sd.shape == (self.T,)
        arr = (TrackResult * self.T)()
        self._chk(lib().pfmpe_track_frame_device_batch(self._h, float(time_to_predict), _d(ip), cn.ctypes.data_as(C.POINTER(C.c_int32)),
                                                       sd.ctypes.data_as(C.POINTER(C.c_uint32)), arr))
        return self._unpack_track(arr)

    def unpack_track(self, arr, tracker, n_frames):
        return self._unpack_track([arr[tracker * n_frames + f] for f in range(n_frames)])

    def track_results(self, first_frame, n_frames, tracker=0):
        arr = (TrackResult * n_frames)()
        self._chk(lib().pfmpe_track_results(self._h, tracker, first_frame, n_frames, arr))
        return self._unpack_track(arr)

    @staticmethod
    def _unpack_track(arr):
        out = []
        for r in arr:
            nc = r.n_correspondences
            out.append(dict(status=r.status, flag_fail=r.flag_fail, iterations=r.iterations,
                            gn_iterations=r.gn_iterations, most_likely_pose_idx=r.most_likely_pose_idx,
                            max_count=r.max_count, highest_prob=r.highest_prob, weight_sum=r.weight_sum,
                            pose=from_colmajor(np.array(r.pose[:])), cov=np.array(r.cov[:]).reshape(6, 6).T.copy(),
                            winner_pose=from_colmajor(np.array(r.winner_pose[:])),
                            correspondences=np.array(r.correspondences[: 2 * nc], dtype=np.uint32).reshape(-1, 2)))
        return out

    # streams / timing
    def set_stream(self, cuda_stream_ptr):
        self._chk(lib().pfmpe_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def sync(self):
        self._chk(lib().pfmpe_sync(self._h))

    def last_launch_ms(self):
        ms = C.c_float()
        self._chk(lib().pfmpe_last_launch_ms(self._h, C.byref(ms)))
        return ms.value

    def launch_count(self):
        n = C.c_uint64()
        self._chk(lib().pfmpe_launch_count(self._h, C.byref(n)))
        return n.value

    def debug_timeline(self, enable=True):
        a = (C.c_uint64 * 16)()
        self._chk(lib().pfmpe_debug_timeline(self._h, int(enable), a))
        return np.array(list(a), dtype=np.uint64)

    def launch_info(self):
        a = (C.c_int32 * 4)()
        self._chk(lib().pfmpe_launch_info(self._h, a))
        return list(a)


# ---- host tail (FP64, no device needed) --------------------------------------------------
def shard_frame(handles, frame_in):
    """pfmpe_shard_frame: one PF update of a sharded filter for this process's shard handles (host logic in C)."""
    arr = (C.c_void_p * len(handles))(*[h._h for h in handles])
    fo = FrameOut()
    handles[0]._chk(lib().pfmpe_shard_frame(arr, len(handles), C.byref(frame_in), C.byref(fo)))
    return unpack_frame_out(fo)


def shard_track_frame_host(handles, time_to_predict, image_points, seed):
    """pfmpe_shard_track_frame_host: predictPose -> sharded PF update -> Gauss-Newton, all in C."""
    arr = (C.c_void_p * len(handles))(*[h._h for h in handles])
    img = _c64(image_points).reshape(-1, 2)
    tr = TrackResult()
    handles[0]._chk(lib().pfmpe_shard_track_frame_host(arr, len(handles), float(time_to_predict), _d(img), img.shape[0],
                                                       int(seed) & 0xFFFFFFFF, C.byref(tr)))
    r = Handle._unpack_track((TrackResult * 1)(tr))[0]
    r["resampled"] = 1 if r["most_likely_pose_idx"] >= 0 else 0
    return r


def optimise_pose(K, markers, img, pairs, pose):
    k = _c64(K).reshape(9)
    m = _c64(markers).reshape(-1, 3)
    im = _c64(img).reshape(-1, 2)
    pr = np.ascontiguousarray(pairs, dtype=np.uint32).reshape(-1, 2)
    p = colmajor(pose).copy()
    cov = np.zeros(36)
    it = C.c_int()
    rc = lib().pfmpe_optimise_pose(_d(k), _d(m), m.shape[0], _d(im), im.shape[0],
                                   pr.ctypes.data_as(C.POINTER(C.c_uint32)), pr.shape[0], _d(p), _d(cov), C.byref(it))
    if rc != OK:
        raise PfmpeError(rc, lib().pfmpe_strerror(rc).decode())
    return from_colmajor(p), cov.reshape(6, 6).T.copy(), it.value


def predict_pose(prev, cur, t_prev, t_cur, t_pred):
    a, b = colmajor(prev), colmajor(cur)
    P, pred = np.zeros(16), np.zeros(16)
    rc = lib().pfmpe_predict_pose(_d(a), _d(b), t_prev, t_cur, t_pred, _d(P), _d(pred))
    if rc != OK:
        raise PfmpeError(rc, lib().pfmpe_strerror(rc).decode())
    return from_colmajor(P), from_colmajor(pred)


def exponential_map(tw):
    t = _c64(tw)
    out = np.zeros(16)
    lib().pfmpe_exponential_map(_d(t), _d(out))
    return from_colmajor(out)


def logarithm_map(T):
    t = colmajor(T)
    out = np.zeros(6)
    lib().pfmpe_logarithm_map(_d(t), _d(out))
    return out
