"""pfmpe_p3p_histogram_static (no reference counterpart): the correspondence
vote with its scratch in the library's file-scope device array instead of a cudaMalloc buffer.  Collected last on
purpose.  Through the C ABI: its histogram and counters must equal those of the existing
entry point, pfmpe_p3p_histogram -- integer votes, compared exactly, the tolerance the other vote tests use between two
device paths -- on two calls in a row (the array is shared and keeps the previous call's contents), with a different
frame in between, and sizes beyond the stated maximum must be rejected with a clear error."""
import numpy as np
import pytest

from pf_monocular_pose_estimator_b200 import capi, scenes as S
from tests.p3p_util import vote_scene

pytestmark = pytest.mark.gpu

KEYS = ("triples", "p3p_solved", "solutions_projected", "solutions_voted")


def _handle(sc):
    h = capi.Handle(1000, n_leds=5, max_blobs=32)
    h.configure(sc.markers, sc.K, S.DEFAULT_PARAMS)
    return h


def test_static_scratch_vote_matches_the_allocating_entry_point_twice_in_a_row():
    sc, img_a, _ = vote_scene(10)          # 15 detections
    _, img_b, _ = vote_scene(25)           # 30 detections: a larger frame in between leaves other contents behind
    h = _handle(sc)
    want_a, st_a = h.p3p_histogram(img_a)
    want_b, st_b = h.p3p_histogram(img_b)
    assert want_a.sum() > 0 and want_b.sum() > 0
    for img, want, st in ((img_a, want_a, st_a), (img_a, want_a, st_a), (img_b, want_b, st_b), (img_a, want_a, st_a)):
        got, sg = h.p3p_histogram_static(img)
        assert got.shape == want.shape and (got == want).all()
        assert all(sg[k] == st[k] for k in KEYS), (sg, st)
    h.close()
    # a second handle (same device) shares the array: still correct when the calls are serialised
    h2 = _handle(sc)
    got, sg = h2.p3p_histogram_static(img_b)
    assert (got == want_b).all() and all(sg[k] == st_b[k] for k in KEYS)
    h2.close()


def test_static_scratch_vote_rejects_sizes_beyond_its_maximum():
    sc, img, _ = vote_scene(3)
    h = _handle(sc)
    rng = np.random.default_rng(1)
    big = np.c_[rng.uniform(0, 752, 257), rng.uniform(0, 480, 257)]
    with pytest.raises(capi.PfmpeError) as e:
        h.p3p_histogram_static(big)
    assert e.value.code == capi.ERR_INVALID and "PFMPE_P3P_STATIC_MAX_POINTS" in str(e.value)
    with pytest.raises(capi.PfmpeError) as e:
        h.p3p_histogram_static(img[:2])
    assert e.value.code == capi.ERR_INVALID
    got, _ = h.p3p_histogram_static(img)            # the handle is still usable afterwards
    assert (got == h.p3p_histogram(img)[0]).all()
    h.close()
