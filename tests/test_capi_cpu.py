"""CPU tests of the C-ABI library: it loads, exports every symbol include/pfmpe.h declares, its
FP64 host tail matches the oracle, and it fails loudly (no CPU fallback) without a CUDA device."""
import os
import re

import numpy as np
import pytest

from oracle import orc
from pf_monocular_pose_estimator_b200 import capi, scenes as S
from tests.util import perturb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_loads_and_exports_every_declared_symbol():
    L = capi.lib()
    hdr = open(os.path.join(ROOT, "include", "pfmpe.h")).read()
    declared = set(re.findall(r"\b(pfmpe_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(capi.EXPORTS), declared ^ set(capi.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert L.pfmpe_version() == 1
    assert L.pfmpe_strerror(0) == b"ok"


def test_no_cpu_fallback_without_device():
    if _has_gpu():
        pytest.skip("a CUDA device is present")
    with pytest.raises(capi.PfmpeError) as e:
        capi.Handle(100)
    assert e.value.code == capi.ERR_NO_DEVICE


def test_create_rejects_bad_config():
    for kw in (dict(n_particles=1), dict(n_particles=100, n_leds=9), dict(n_particles=100, max_blobs=0)):
        with pytest.raises(capi.PfmpeError) as e:
            capi.Handle(**kw)
        assert e.value.code == capi.ERR_INVALID


def test_lie_maps_match_oracle():
    rng = np.random.default_rng(11)
    for _ in range(50):
        tw = rng.uniform(-0.6, 0.6, 6)
        T = capi.exponential_map(tw)
        np.testing.assert_allclose(T, orc.exponential_map(tw), atol=1e-15)
        np.testing.assert_allclose(capi.logarithm_map(T), orc.logarithm_map(T), atol=1e-13)
    np.testing.assert_array_equal(capi.exponential_map(np.zeros(6)), np.eye(4))
    np.testing.assert_array_equal(capi.logarithm_map(np.eye(4)), np.zeros(6))


def test_predict_pose_matches_oracle():
    sc = S.Scene()
    for f in (2, 10, 57):
        a, b = sc.pose(f - 2), sc.pose(f - 1)
        P, pred = capi.predict_pose(a, b, sc.time(f - 2), sc.time(f - 1), sc.time(f))
        Po, predo = orc.predict_pose(a, b, sc.time(f - 2), sc.time(f - 1), sc.time(f))
        np.testing.assert_allclose(P, Po, atol=1e-13)
        np.testing.assert_allclose(pred, predo, atol=1e-13)
    # non-orthonormal rotation block (FP32 particle): the true inverse is used, as Eigen's .inverse()
    a32 = a.astype(np.float32).astype(np.float64)
    P, _ = capi.predict_pose(a32, b, 0.0, 0.02, 0.04)
    Po, _ = orc.predict_pose(a32, b, 0.0, 0.02, 0.04)
    np.testing.assert_allclose(P, Po, atol=1e-13)


def test_gauss_newton_matches_oracle():
    rng = np.random.default_rng(4)
    for trial in range(10):
        T = S.true_pose(0.3 * trial)
        uv = S.project(S.K_README, T, S.MARKERS_5) + rng.normal(0, 0.3, (5, 2))
        img = np.concatenate([uv, rng.uniform(0, 400, (4, 2))])
        perm = rng.permutation(9)
        img = img[perm]
        inv = np.argsort(perm)
        keep = 5 if trial % 3 else 4
        pairs = np.array([[j + 1, inv[j] + 1] for j in range(keep)], dtype=np.uint32)
        start = perturb(T, rng, 0.01, 0.01).astype(np.float32).astype(np.float64)
        pose, cov, it = capi.optimise_pose(S.K_README, S.MARKERS_5, img, pairs, start)
        pose_o, cov_o, it_o = orc.optimise_pose(S.K_README, S.MARKERS_5, img, pairs, start)
        np.testing.assert_allclose(pose, pose_o, atol=1e-9)
        np.testing.assert_allclose(cov, cov_o, rtol=1e-6, atol=1e-12)
        assert abs(it - it_o) <= 2


def test_optimise_pose_rejects_bad_correspondences():
    T = S.true_pose(0.1)
    uv = S.project(S.K_README, T, S.MARKERS_5)
    with pytest.raises(capi.PfmpeError):
        capi.optimise_pose(S.K_README, S.MARKERS_5, uv, np.array([[6, 1]], dtype=np.uint32), T)
    with pytest.raises(capi.PfmpeError):
        capi.optimise_pose(S.K_README, S.MARKERS_5, uv, np.array([[1, 9]], dtype=np.uint32), T)
