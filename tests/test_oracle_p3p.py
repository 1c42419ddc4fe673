"""CPU tests of the P3P / initialisation-vote oracle (oracle/p3p_oracle.cpp).  The reference
has no tests for this path, so the restatement is pinned through properties: a P3P solution reproduces the
pose the bearings were generated from, every real solution reprojects its three points exactly, the quartic
solver returns the roots of polynomials built from known roots, and the vote finds the true correspondences."""
import json
import os
from math import comb

import numpy as np
import pytest

from oracle import orc
from pf_monocular_pose_estimator_b200 import scenes as S
from tests.p3p_util import random_problems, vote_scene

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "p3p_votes.json")


def test_quartic_known_real_roots():
    rng = np.random.default_rng(5)
    for _ in range(200):
        r = np.sort(rng.uniform(-1, 1, 4))
        a = rng.uniform(0.5, 2.0) * rng.choice([-1, 1])
        got = np.sort(orc.p3p_solve_quartic(a * np.poly(r)))
        np.testing.assert_allclose(got, r, atol=2e-6)   # Ferrari's closed form loses digits when roots are close


def test_quartic_matches_numpy_roots_real_parts():
    rng = np.random.default_rng(6)
    for _ in range(200):
        f = rng.uniform(-1, 1, 5)
        f[0] = rng.uniform(0.2, 1.0) * rng.choice([-1, 1])
        got = np.sort(orc.p3p_solve_quartic(f))
        want = np.sort(np.roots(f).real)
        np.testing.assert_allclose(got, want, atol=1e-6, rtol=1e-6)


def test_p3p_recovers_the_generating_pose():
    fv, wp, poses = random_problems(300, seed=3, wrong_fraction=0.0)
    for i in range(fv.shape[0]):
        rc, sol = orc.p3p_compute_poses(fv[i], wp[i])
        assert rc == 0
        best = np.inf
        for s in sol:
            if not np.isfinite(s).all():
                continue
            H = np.eye(4); H[:3, :4] = s
            best = min(best, np.abs(np.linalg.inv(H) - poses[i]).max())
        assert best < 1e-7, (i, best)


def test_p3p_real_solutions_reproject_their_points():
    fv, wp, _ = random_problems(300, seed=4, wrong_fraction=1.0)
    checked = 0
    for i in range(fv.shape[0]):
        rc, sol = orc.p3p_compute_poses(fv[i], wp[i])
        assert rc == 0
        for s in sol:
            if not np.isfinite(s).all():
                continue
            R, Cc = s[:, :3], s[:, 3]
            if abs(np.linalg.det(R) - 1) > 1e-9:
                continue
            pc = (wp[i].T - Cc) @ R          # R^T (X - C) for the three points (rows)
            b = pc / np.linalg.norm(pc, axis=1, keepdims=True)
            err = np.abs(b - fv[i].T).max()
            if err < 1e-6:
                checked += 1
    assert checked > 100   # real roots give exact poses; real parts of complex roots do not and are not counted


def test_p3p_colinear_world_points_are_rejected():
    fv, _, _ = random_problems(1, seed=1)
    wp = np.array([[0.0, 0.1, 0.2], [0, 0, 0], [0, 0, 0]])
    rc, _ = orc.p3p_compute_poses(fv[0], wp)
    assert rc == -1


def test_image_vectors_are_unit_bearings():
    img = S.Scene().blobs(3)
    v = orc.image_vectors(S.K_README, img)
    np.testing.assert_allclose(np.linalg.norm(v, axis=1), 1.0, atol=1e-15)
    np.testing.assert_allclose(v[:, 0] / v[:, 2], (img[:, 0] - S.K_README[0, 2]) / S.K_README[0, 0], rtol=1e-14)


@pytest.mark.parametrize("n_out", [0, 3, 10])
def test_vote_counters_and_true_correspondences(n_out):
    sc, img, true_blob = vote_scene(n_out)
    hist, st = orc.p3p_histogram(sc.K, sc.markers, img, 5.0)
    D, L = img.shape[0], sc.markers.shape[0]
    assert st[0] == comb(D, 3) and st[1] == comb(D, 3) * L * (L - 1) * (L - 2)
    assert st[3] <= st[2] <= 4 * st[1]
    # every voting solution casts its three own pairs plus at least one more
    assert hist.sum() >= 4 * st[3]
    if n_out == 0:
        assert list(hist.argmax(0)) == true_blob
    if n_out <= 3:   # among the true blobs the true pairing dominates every LED column (outlier-heavy scenes need the
        sub = hist[true_blob]   # normalisation of correspondencesFromHistogram, which is the caller's)
        assert (sub.argmax(0) == np.arange(L)).all()


def test_vote_needs_five_detections():
    sc, img, _ = vote_scene(0)
    for D in (3, 4):
        hist, st = orc.p3p_histogram(sc.K, sc.markers, img[:D], 5.0)
        assert hist.sum() == 0 and st[0] == 0   # cd < 5 (:1578) filters every triple


def test_vote_four_leds():
    sc, img, true_blob = vote_scene(2, markers=S.MARKERS_4)
    hist, st = orc.p3p_histogram(sc.K, sc.markers, img, 5.0)
    assert hist.shape == (6, 4) and st[1] == comb(6, 3) * 24
    assert (hist[true_blob].argmax(0) == np.arange(4)).all()


def test_vote_golden_fixture():
    """Regression fixture written by tests/golden/make_p3p_golden.py from this oracle (the reference itself cannot
    be built here: parity unpinned)."""
    g = json.load(open(GOLDEN))
    for case in g["cases"]:
        sc, img, _ = vote_scene(case["n_outliers"], frame=case["frame"])
        np.testing.assert_array_equal(img, np.array(case["image_points"]))
        hist, st = orc.p3p_histogram(sc.K, sc.markers, img, g["tol"])
        assert hist.tolist() == case["hist"] and list(st) == case["stats"]


def test_combinations_follow_the_reference_order():
    """Combinations::combinationsNoReplacement restated literally (combinations.cpp:60-128) = lexicographic order."""
    from itertools import combinations
    for n in (3, 4, 5, 6, 8):
        assert orc.combinations(n, 3).tolist() == [list(c) for c in combinations(range(1, n + 1), 3)]
    assert orc.combinations(5, 1).reshape(-1).tolist() == [1, 2, 3, 4, 5]


def test_check_correspondences_accepts_the_true_pairing_only():
    sc, img, true_blob = vote_scene(10)
    corr = np.array([[j + 1, true_blob[j] + 1] for j in range(5)])
    r = orc.check_correspondences(sc.K, sc.markers, img, corr, 5.0, 0.75)
    assert len(r) == 10 and all(x["rc"] == 0 and x["valid"] == 1 and 0 <= x["best"] < 4 for x in r)
    dev = [np.abs(x["pose"] - sc.pose(7)).max() for x in r]   # the selected solution is the object's pose, up to what
    assert np.median(dev) < 0.03 and max(dev) < 0.2           # 0.3 px of blob noise does to a three-point pose
    assert all(x["squared_error"] <= 2 * 5.0 ** 2 for x in r)
    swapped = corr.copy()
    swapped[[0, 1], 1] = swapped[[1, 0], 1]
    assert not any(x["valid"] for x in orc.check_correspondences(sc.K, sc.markers, img, swapped, 5.0, 0.75))
    # three pairs: nothing is left to verify with, the certainty is 0/0 and never reaches a threshold (:1131, :1412)
    r3 = orc.check_correspondences(sc.K, sc.markers, img, corr[:3], 5.0, 0.75)
    assert len(r3) == 1 and r3[0]["valid"] == 0
    # four pairs, certainty threshold 1: the single unused pair must reproject within the tolerance
    r4 = orc.check_correspondences(sc.K, sc.markers, img, corr[:4], 5.0, 1.0)
    assert len(r4) == 4 and all(x["valid"] for x in r4)
