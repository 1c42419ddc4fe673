"""One filter sharded over several handles (BASELINE configs[4]): the sharded filter must reproduce the
unsharded CUDA filter bit for bit (weights, propagated particles, resampled set, winner, record),
because RNG addressing, strata and tie rules use global numbers.  Logical shards on one GPU
(Loopback communicator); the NCCL path is run by tools/shard_nccl_check.py under torchrun."""
import numpy as np
import pytest

from pf_monocular_pose_estimator_b200 import capi, scenes, sharded

from .cuda_util import make_handle
from .host_loop import HostTracker
from .util import frame_inputs

pytestmark = pytest.mark.gpu

FIELDS = ["status", "iterations", "best_iteration", "resampled", "most_likely_particle_idx", "most_likely_pose_idx",
          "max_count"]


def _make_sharded(fi, N, W, **kw):
    D = np.asarray(fi["img"]).reshape(-1, 2).shape[0]
    s = sharded.ShardedFilter(N, sharded.Loopback(W), n_leds=fi["markers"].shape[0], max_blobs=max(D, 1), **kw)
    s.configure(fi["markers"], fi["K"], fi["params"])
    return s


def _same_frame(a, b, ha, hb):
    for k in FIELDS:
        assert a[k] == b[k], k
    assert a["highest_prob"] == b["highest_prob"] and a["weight_sum"] == b["weight_sum"]
    assert (a["winner_pose"] == b["winner_pose"]).all()
    assert (a["correspondences"] == b["correspondences"]).all()
    assert (ha.weights(False) == hb.weights(False)).all()
    assert (ha.particles(0) == hb.particles(0)).all()
    if a["resampled"]:
        assert (ha.particles(1) == hb.particles(1)).all()


@pytest.mark.parametrize("N,W,kw", [(4096, 8, {}), (8192, 4, dict(rebalance="exchange")), (6144, 2, dict(force_mode=2, grid_cap=3)),
                                    (200000, 2, {}), (200000, 2, dict(rebalance="exchange")), (40960, 5, {}),
                                    (8192, 4, dict(mail=False))])
def test_logical_shards_match_unsharded_filter(N, W, kw):
    fi = frame_inputs(N, frame=7)
    ref = make_handle(fi, N)
    sh = _make_sharded(fi, N, W, **kw)
    assert sh.partial_exchange == ("mail" if kw.get("mail", True) and kw.get("rebalance") != "exchange" else "all_gather")
    ref.set_particles(fi["E"])
    sh.set_particles(fi["E"])
    args = (fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], fi["isi"], fi["dt"])
    for f in range(3):                    # later frames start from the rebalanced resampled set
        a = ref.update(*args, fi["seed"] + f)
        b = sh.update(*args, fi["seed"] + f)
        assert a["resampled"] == 1
        _same_frame(a, b, ref, sh)
    ref.close()
    sh.close()


@pytest.mark.parametrize("rebalance", ["p2p", "exchange"])
def test_sharded_dominant_sources_and_empty_ranks(rebalance):
    """A handful of sources (all on shard 0) own every output slot: the other shards produce nothing and
    receive their whole particle set through the rebalance; then a single dominant source."""
    N, W = 32768, 4
    fi = frame_inputs(N, frame=14)
    far = fi["E"][40:].copy()
    far[:, 0, 3] += 5.0
    fi["E"][40:] = far
    ref = make_handle(fi, N)
    sh = _make_sharded(fi, N, W, rebalance=rebalance)
    assert sh.rebalance == rebalance
    args = (fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], fi["isi"], fi["dt"], fi["seed"])
    for step in range(2):
        ref.set_particles(fi["E"])
        sh.set_particles(fi["E"])
        a, b = ref.update(*args), sh.update(*args)
        assert a["resampled"] == 1 and a["max_count"] > (500 if step == 0 else N // 3)
        _same_frame(a, b, ref, sh)
        fi["E"][2:] = far[0]
        # sources on the LAST shard only
        E2 = fi["E"][::-1].copy()
        ref.set_particles(E2)
        sh.set_particles(E2)
        a, b = ref.update(*args), sh.update(*args)
        _same_frame(a, b, ref, sh)
    ref.close()
    sh.close()


def test_sharded_not_accepted_and_first_frame():
    N, W = 8192, 4
    # no blob near any projection: not accepted (status 4), argmax/argmin winner, no resampling
    fi = frame_inputs(N, frame=9)
    fi["img"] = np.asarray(fi["img"]) * 0 + np.array([5.0, 5.0])
    ref = make_handle(fi, N)
    sh = _make_sharded(fi, N, W)
    ref.set_particles(fi["E"])
    sh.set_particles(fi["E"])
    args = [fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], fi["isi"], fi["dt"], fi["seed"]]
    a, b = ref.update(*args), sh.update(*args)
    assert a["status"] == 4 and a["resampled"] == 0 and a["iterations"] == 80
    _same_frame(a, b, ref, sh)
    ref.close()
    sh.close()
    # it_since_initialized == 0: accepted but not resampled (:666)
    fi = frame_inputs(N, frame=9, isi=0)
    ref = make_handle(fi, N)
    sh = _make_sharded(fi, N, W)
    ref.set_particles(fi["E"])
    sh.set_particles(fi["E"])
    args = [fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], 0, fi["dt"], fi["seed"]]
    a, b = ref.update(*args), sh.update(*args)
    assert a["status"] == 1 and a["resampled"] == 0
    _same_frame(a, b, ref, sh)
    ref.close()
    sh.close()


def test_sharded_tracking_loop_matches_unsharded():
    """The reference-side loop (predict -> PF -> GN) runs unchanged on top of ShardedFilter.update."""
    N, W = 16384, 4
    sc = scenes.Scene(0)
    ref = capi.Handle(N, n_trackers=1, n_leds=5, max_blobs=sc.blobs(0).shape[0] + 4)
    ref.configure(sc.markers, sc.K, sc.params)
    sh = sharded.ShardedFilter(N, sharded.Loopback(W), n_leds=5, max_blobs=sc.blobs(0).shape[0] + 4)
    sh.configure(sc.markers, sc.K, sc.params)
    ta, tb = HostTracker(ref, sc), HostTracker(sh, sc)
    for f in range(2, 14):
        a, b = ta.frame(f), tb.frame(f)
        for k in FIELDS:
            assert a[k] == b[k], (f, k)
        assert (a["pose"] == b["pose"]).all() and (a["cov"] == b["cov"]).all()
    assert (ref.particles(1) == sh.particles(1)).all()
    ref.close()
    sh.close()


@pytest.mark.parametrize("N,W", [(8192, 4), (200000, 2)])
def test_sharded_weighted_mean_and_covariance(N, W):
    """K7 under particle sharding (an exchange of 12 + 21 doubles): the weighted mean pose and covariance the
    sharded C frame loop returns against the unsharded filter's (same particles and weights bit for bit; the sums are
    taken in a different order, hence 1e-10) -- which tests/test_cuda_parity.py checks against the oracle."""
    fi = frame_inputs(N, frame=9)
    ref = make_handle(fi, N, flags=capi.FLAG_MEAN_COV)
    sh = _make_sharded(fi, N, W, flags=capi.FLAG_MEAN_COV)
    assert sh.loop == "c"
    ref.set_particles(fi["E"])
    sh.set_particles(fi["E"])
    args = (fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], fi["isi"], fi["dt"])
    for f in range(3):
        a = ref.update(*args, fi["seed"] + f)
        b = sh.update(*args, fi["seed"] + f)
        _same_frame(a, b, ref, sh)
        assert np.abs(a["mean_pose"][:3, :3] @ a["mean_pose"][:3, :3].T - np.eye(3)).max() < 1e-12
        np.testing.assert_allclose(b["mean_pose"], a["mean_pose"], rtol=0, atol=1e-10)
        np.testing.assert_allclose(b["mean_cov"], a["mean_cov"], rtol=1e-8, atol=1e-16)
        assert np.abs(a["mean_cov"]).max() > 0
    ref.close()
    sh.close()
