"""include/pfmpe_pose_estimator.hpp -- the C++ mirror of the reference's PoseEstimator interface -- compiled with
g++ against libpfmpe.so and driven like MPENode drives the reference.  CPU: it builds, and without a CUDA device
every frame fails loudly (Flag_Fail 100 = kFlagFailNoDevice, no CPU fallback).  GPU: its poses equal the ctypes path bit for bit."""
import os
import subprocess

import numpy as np
import pytest

from pf_monocular_pose_estimator_b200 import capi, scenes as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pf_monocular_pose_estimator_b200")
N, F, FIRST = 2000, 12, 2


def _build(tmp_path):
    capi.lib()
    exe = str(tmp_path / "pose_estimator_demo")
    subprocess.check_call(["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "pose_estimator_demo.cpp"), "-o", exe,
                           "-L", PKG, "-lpfmpe", "-Wl,-rpath," + PKG])
    return exe


def _sequence(tmp_path):
    sc = S.Scene()
    p = tmp_path / "sequence.txt"
    col = lambda T: " ".join(repr(float(v)) for v in np.asarray(T).T.reshape(-1))      # column-major
    with open(p, "w") as f:
        f.write("%d %d %d %d\n" % (N, F, sc.markers.shape[0], S.pf_seed(FIRST)))
        for m in sc.markers:
            f.write("%r %r %r\n" % tuple(float(v) for v in m))
        f.write(" ".join(repr(float(v)) for v in sc.K.reshape(-1)) + "\n")
        f.write(col(sc.pose(0)) + " %r\n" % sc.time(0))
        f.write(col(sc.pose(1)) + " %r\n" % sc.time(1))
        for k in range(FIRST, FIRST + F):
            b = sc.blobs(k)
            f.write("%r %d\n" % (sc.time(k), b.shape[0]))
            for q in b:
                f.write("%r %r\n" % (float(q[0]), float(q[1])))
    return sc, str(p)


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_cpp_mirror_builds_and_fails_loudly_without_device(tmp_path):
    exe = _build(tmp_path)
    if _has_gpu():
        pytest.skip("a CUDA device is present")
    _, seq = _sequence(tmp_path)
    out = subprocess.run([exe, seq], capture_output=True, text=True, check=True).stdout.splitlines()
    frames = [l for l in out if l[:2] in ("0 ", "1 ")]
    assert len(frames) == F and all(l.startswith("0 100.0") for l in frames)     # found = false, Flag_Fail = kFlagFailNoDevice
    assert any("no usable CUDA device" in l or "no device" in l for l in out)


@pytest.mark.gpu
def test_cpp_mirror_matches_ctypes_path(tmp_path):
    exe = _build(tmp_path)
    sc, seq = _sequence(tmp_path)
    out = subprocess.run([exe, seq], capture_output=True, text=True, check=True).stdout.splitlines()
    frames = [l.split() for l in out if l[:2] in ("0 ", "1 ")]
    assert len(frames) == F
    h = capi.Handle(N, n_leds=5, max_blobs=256)
    h.configure(sc.markers, sc.K, sc.params)
    h.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
    for i, k in enumerate(range(FIRST, FIRST + F)):
        r = h.track_frame_host(sc.time(k), sc.blobs(k), sc.seed(k))
        got = np.array([float(v) for v in frames[i][3:19]]).reshape(4, 4).T
        assert int(frames[i][0]) == 1 and float(frames[i][1]) == r["flag_fail"] and int(frames[i][2]) == r["gn_iterations"]
        np.testing.assert_array_equal(got, r["pose"])
        assert np.abs(got[:3, 3] - sc.pose(k)[:3, 3]).max() < 0.02           # follows the true trajectory
    assert out[-1] == "particles %d resampled %d weights %d correspondences 5" % (N, N, N)
    h.close()
