"""The initialisation side: the device stand-ins of the C++ mirror (correspondenceHistogram,
checkCorrespondenceSets, seedInitialisedObject) strung together by tests/pose_estimator_init.cpp with the reference's
host tail of initialise(), which is restated in oracle/initialise_host.hpp (test infrastructure).  CPU: the oracle's
host functions -- correspondencesFromHistogram (:1134-1288), computeTransformation (:2139-2161) -- against independent
NumPy restatements.  GPU: the initialisation of a synthetic frame finds the true LED / detection pairing and the object's
pose, and the particle filter tracks on from the seeded particles (one object, and two objects at one timestamp)."""
import os
import subprocess

import numpy as np
import pytest

from pf_monocular_pose_estimator_b200 import capi, scenes as S
from tests.p3p_util import vote_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pf_monocular_pose_estimator_b200")


def _build(tmp_path):
    capi.lib()
    exe = str(tmp_path / "pose_estimator_init")
    subprocess.check_call(["g++", "-std=c++11", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "pose_estimator_init.cpp"), "-o", exe,
                           "-L", PKG, "-lpfmpe", "-Wl,-rpath," + PKG])
    return exe


def correspondences_from_histogram_np(hist):
    """NumPy restatement of pose_estimator.cpp:1134-1288 with bInitialisation = true."""
    R, C = hist.shape
    thr = 1.3 / (R * C)
    hp = hist.astype(np.float64)
    row = hist.sum(1).astype(np.uint32)
    for c in range(C):
        cs = np.uint32(hist[:, c].sum())
        if cs == 0:
            continue
        for r in range(R):
            den = float(np.uint32(cs * row[r]))                # unsigned product, :1155
            v = hp[r, c] ** 2 / den if den != 0 else float("nan")
            v = 0.0 if not (v > 0.0) else v                    # std::max(0.0, NaN) = 0.0
            hp[r, c] = v if v >= thr else 0.0
    u_prob = [[hp[b, a] for b in range(R) if hp[b, a] != 0] for a in range(C)]
    u_num = [[b + 1 for b in range(R) if hp[b, a] != 0] for a in range(C)]
    n_v = [len(x) for x in u_num]
    N = int(np.prod([max(1, k) for k in n_v]))
    combs, probs = [], []
    for i in range(N):
        prob, n, comb = 1.0, 1, []
        for led in range(C - 1, -1, -1):
            if n_v[led] > 0:
                d = (i // n) % n_v[led]
                prob *= u_prob[led][d]
                comb.append(u_num[led][d])
                n *= max(1, n_v[led])
            else:
                comb.append(0)
        combs.append(comb[::-1]); probs.append(prob)
    probs = np.array(probs) / np.sum(probs)
    out = []
    for _ in range(N):
        k = int(np.argmax(probs))
        probs[k] = 0
        det = combs[k]
        if len(set(det)) < len(det):                           # checkAmbiguity
            continue
        out.append([(led + 1, det[led]) for led in range(C) if det[led] != 0])
    return out


def test_correspondences_from_histogram_matches_numpy(tmp_path):
    from oracle import orc
    exe = _build(tmp_path)
    cases = []
    for n_out in (0, 3, 10):
        sc, img, _ = vote_scene(n_out)
        hist, _ = orc.p3p_histogram(sc.K, sc.markers, img, 5.0)
        cases.append(hist)
    rng = np.random.default_rng(3)
    cases.append(rng.integers(0, 4, (6, 5)).astype(np.uint32) * rng.integers(0, 2, (6, 5)).astype(np.uint32))
    z = np.zeros((7, 5), dtype=np.uint32); z[2, 0] = 5; z[3, 1] = 4; z[3, 2] = 4
    cases.append(z)                                            # a missing LED column and an ambiguous detection
    for i, hist in enumerate(cases):
        p = tmp_path / ("hist%d.txt" % i)
        with open(p, "w") as f:
            f.write("%d %d\n" % hist.shape + " ".join(str(int(v)) for v in hist.reshape(-1)) + "\n")
        out = subprocess.run([exe, "hist", str(p)], capture_output=True, text=True, check=True).stdout.splitlines()
        got = [[tuple(int(x) for x in tok.split(":")) for tok in line.split()] for line in out]
        assert got == correspondences_from_histogram_np(hist), i
    # the true pairing of a clean scene comes first
    sc, img, true_blob = vote_scene(3)
    want = [(j + 1, true_blob[j] + 1) for j in range(5)]
    assert correspondences_from_histogram_np(orc.p3p_histogram(sc.K, sc.markers, img, 5.0)[0])[0] == want


def test_compute_transformation_matches_svd(tmp_path):
    exe = _build(tmp_path)
    rng = np.random.default_rng(8)
    for i in range(6):
        T = S.true_pose(rng.uniform(0, 30), rng.uniform(0, 5))
        obj = S.MARKERS_5 if i < 4 else np.c_[rng.uniform(-0.2, 0.2, (5, 2)), np.zeros(5)]    # i >= 4: coplanar (rank 2)
        rep = obj @ T[:3, :3].T + T[:3, 3] + (rng.normal(0, 1e-3, (5, 3)) if i % 2 else 0.0)
        p = tmp_path / ("x%d.txt" % i)
        with open(p, "w") as f:
            f.write("5\n" + "\n".join("%r %r %r" % tuple(map(float, r)) for r in obj) + "\n" +
                    "\n".join("%r %r %r" % tuple(map(float, r)) for r in rep) + "\n")
        got = np.array([float(v) for v in subprocess.run([exe, "xform", str(p)], capture_output=True, text=True,
                                                         check=True).stdout.split()]).reshape(4, 4).T
        ob, rb = obj - obj.mean(0), rep - rep.mean(0)
        U, _, Vt = np.linalg.svd(ob.T @ rb)                      # :2147-2153: R = V U^T
        R = Vt.T @ U.T
        if np.linalg.det(R) < 0:                                # rank-2 input: the mirror completes to a proper rotation
            Vt[2] *= -1
            R = Vt.T @ U.T
        want = np.eye(4); want[:3, :3] = R; want[:3, 3] = rep.mean(0) - R @ obj.mean(0)
        np.testing.assert_allclose(got, want, atol=1e-9)
        assert abs(np.linalg.det(got[:3, :3]) - 1) < 1e-9
        if i % 2 == 0:
            np.testing.assert_allclose(got, T, atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("n_out,n_obj", [(0, 1), (5, 1), (10, 1), (3, 2)])
def test_initialise_then_track(tmp_path, n_out, n_obj):
    exe = _build(tmp_path)
    sc = S.Scene(n_outliers=n_out)
    f0, F, N = 7, 10, 2000
    p = tmp_path / "init.txt"
    with open(p, "w") as f:
        f.write("%d %d %d %d\n" % (N, 5, S.pf_seed(f0 + 1), n_obj))
        for m in sc.markers:
            f.write("%r %r %r\n" % tuple(float(v) for v in m))
        f.write(" ".join(repr(float(v)) for v in sc.K.reshape(-1)) + "\n")
        for k in range(f0, f0 + F + 1):
            if k == f0 + 1:
                f.write("%d\n" % F)
            b = sc.blobs(k)
            f.write("%r %d\n" % (sc.time(k), b.shape[0]) + "\n".join("%r %r" % (float(q[0]), float(q[1])) for q in b) + "\n")
    out = subprocess.run([exe, "init", str(p)], capture_output=True, text=True, check=True).stdout.splitlines()
    head = out[0].split()
    assert head[:4] == ["init", "1", "flag", "0.0"], out[:3]
    img = sc.blobs(f0)
    uv = S.project(sc.K, sc.pose(f0), sc.markers)
    true_blob = [int(np.argmin(((img - u) ** 2).sum(1))) for u in uv]
    assert head[4:] == ["%d:%d" % (j + 1, true_blob[j] + 1) for j in range(5)]          # the true pairing
    pose0 = np.array([float(v) for v in out[1].split()[1:]]).reshape(4, 4).T
    assert np.abs(pose0 - sc.pose(f0)).max() < 0.03                                      # Gauss-Newton on 5 pairs, 0.3 px noise
    frames = out[2:]
    assert len(frames) == F
    for i, line in enumerate(frames):
        tok = line.split()
        assert tok[0] == "1", (i, line)                                                  # the PF tracks on from the seeds
        pose = np.array([float(v) for v in tok[2:18]]).reshape(4, 4).T
        assert np.abs(pose[:3, 3] - sc.pose(f0 + 1 + i)[:3, 3]).max() < 0.03
        last = np.array([float(v) for v in tok[18:34]]).reshape(4, 4).T                   # the last object (same detections)
        assert np.isfinite(last).all() and np.abs(last[:3, 3] - sc.pose(f0 + 1 + i)[:3, 3]).max() < 0.03
