"""Loader of tests/_hostmath.so: the device math headers compiled for the host (TEST ARTEFACT ONLY)."""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle import orc

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_hostmath.so")
_SRC = os.path.join(_HERE, "hostmath.cpp")
_CSRC = os.path.join(os.path.dirname(_HERE), "pf_monocular_pose_estimator_b200", "csrc")
_lib = None


def lib():
    global _lib
    if _lib is None:
        deps = [_SRC] + [os.path.join(_CSRC, f) for f in ("pf_math.cuh", "pf_tail.cuh")]
        if not os.path.exists(_SO) or any(os.path.getmtime(d) > os.path.getmtime(_SO) for d in deps):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared",
                                   "-Wno-unknown-pragmas", "-o", _SO, _SRC])
        L = C.CDLL(_SO)
        L.hm_rng_skip.restype = C.c_uint32
        L.hm_rng_skip.argtypes = [C.c_uint32, C.c_uint64]
        L.hm_canonical_f64.restype = C.c_double
        L.hm_canonical_f64.argtypes = [C.c_uint32, C.c_uint32]
        L.hm_canonical_f32.restype = C.c_float
        L.hm_canonical_f32.argtypes = [C.c_uint32, C.c_uint32]
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def frame(fi):
    """Replay one frame of the kernel algorithm on the host.  fi: dict from tests.util.frame_inputs."""
    prm = fi["params"]
    markers = np.ascontiguousarray(fi["markers"], dtype=np.float64)
    L = markers.shape[0]
    img = np.ascontiguousarray(fi["img"], dtype=np.float64).reshape(-1, 2)
    D = img.shape[0]
    E = orc.colmajor(fi["E"])
    N = E.shape[0]
    K = np.ascontiguousarray(fi["K"], dtype=np.float64).reshape(9)
    cur, pred, P, Cm = (orc.colmajor(fi[k]) for k in ("T_cur", "T_pred", "P", "Cm"))
    pr = np.array([prm["minAngularNoise"], prm["maxAngularNoise"], prm["minTransitionNoise"], prm["maxTransitionNoise"],
                   prm["back_projection_pixel_tolerance"], prm["back_projection_pixel_tolerance_PF"]])
    dg = np.array(prm.get("bMarkerDowngrade", [0] * 5), dtype=np.int32)
    T = np.zeros((N, 12), dtype=np.float32); w = np.zeros(N, dtype=np.float32)
    kend = np.zeros(N, dtype=np.int32); idx = np.zeros(N, dtype=np.int32)
    sc = np.zeros(16, dtype=np.int32); sfx = C.c_longlong(0); pairs = np.zeros(16, dtype=np.uint32)
    d = C.c_double
    rc = lib().hm_frame(N, L, D, _p(markers, d), _p(K, d), _p(img, d), _p(E, d), _p(cur, d), _p(pred, d), _p(P, d),
                        _p(Cm, d), C.c_uint(fi["isi"]), C.c_double(fi["dt"]), C.c_uint32(fi["seed"] & 0xFFFFFFFF),
                        _p(pr, d), _p(dg, C.c_int), _p(T, C.c_float), _p(w, C.c_float), _p(kend, C.c_int),
                        _p(idx, C.c_int), _p(sc, C.c_int), C.byref(sfx), _p(pairs, C.c_uint))
    assert rc == 0
    T4 = np.zeros((N, 4, 4)); T4[:, :3, :] = T.reshape(N, 3, 4); T4[:, 3, 3] = 1
    counts = np.diff(np.concatenate([[0], kend])).astype(np.uint32)
    return dict(T=T4, w_raw=w.astype(np.float64), kend=kend, idx=idx, count=counts, iterations=int(sc[0]),
                best_iteration=int(sc[1]), accepted=int(sc[2]), resampled=int(sc[3]),
                most_likely_particle_idx=int(sc[4]), winner=int(sc[5]), max_count=int(sc[6]),
                pairs=pairs[: 2 * sc[7]].reshape(-1, 2).copy(), sum_fx=sfx.value,
                highest_prob=float(sc[8:9].view(np.float32)[0]))


def resample(w32, seed, iters):
    w32 = np.ascontiguousarray(w32, dtype=np.float32)
    N = w32.shape[0]
    kend = np.zeros(N, dtype=np.int32); idx = np.zeros(N, dtype=np.int32); sfx = C.c_longlong(0)
    rc = lib().hm_resample(N, _p(w32, C.c_float), C.c_uint32(seed & 0xFFFFFFFF), iters, _p(kend, C.c_int),
                           _p(idx, C.c_int), C.byref(sfx))
    counts = np.diff(np.concatenate([[0], kend])).astype(np.uint32)
    return rc, idx, counts, sfx.value
