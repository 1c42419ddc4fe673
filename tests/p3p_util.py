"""Shared inputs for the P3P / initialisation-vote tests."""
import numpy as np

from pf_monocular_pose_estimator_b200 import scenes as S


def random_problems(n, seed=3, markers=S.MARKERS_5, K=S.K_README, wrong_fraction=0.5):
    """n P3P problems: bearing vectors of three projected LEDs at a random pose; for a fraction of them the
    LED triple handed to the solver is NOT the one that was projected (the initialisation tries every
    permutation, so most of its solves are of this kind).  Returns fv (n,3,3), wp (n,3,3) with the vectors /
    points as COLUMNS, and the true poses."""
    rng = np.random.default_rng(seed)
    L = markers.shape[0]
    fv = np.zeros((n, 3, 3)); wp = np.zeros((n, 3, 3)); poses = np.zeros((n, 4, 4))
    for i in range(n):
        T = S.true_pose(rng.uniform(0, 60.0), rng.uniform(0, 6.0))
        tri = rng.permutation(L)[:3]
        uv = S.project(K, T, markers[tri])
        v = np.stack([(uv[:, 0] - K[0, 2]) / K[0, 0], (uv[:, 1] - K[1, 2]) / K[1, 1], np.ones(3)], axis=1)
        v /= np.linalg.norm(v, axis=1, keepdims=True)
        given = tri if rng.uniform() > wrong_fraction else rng.permutation(L)[:3]
        fv[i] = v.T
        wp[i] = markers[given].T
        poses[i] = T
    return fv, wp, poses


def vote_scene(n_outliers, frame=7, tracker=0, markers=S.MARKERS_5):
    sc = S.Scene(tracker=tracker, n_outliers=n_outliers, markers=markers)
    img = sc.blobs(frame)
    uv = S.project(sc.K, sc.pose(frame), sc.markers)
    true_blob = [int(np.argmin(((img - u) ** 2).sum(1))) for u in uv]
    return sc, img, true_blob
