"""Parity at the sizes BASELINE.json names (missing #2 / #3): configs[2] against the oracle at full size,
configs[3] as the real 4096 x 4096 batch with randomly chosen trackers against the oracle, configs[4] (2^26 particles)
sharded against unsharded by device-side checksums, and the end-to-end index agreement report of SURVEY 7.3(b)."""
import numpy as np
import pytest

from oracle import orc
from pf_monocular_pose_estimator_b200 import capi, scenes as S, sharded

from . import parity
from .cuda_util import collect, make_handle, run_frame
from .util import frame_inputs

pytestmark = pytest.mark.gpu


def _oracle(fi):
    return orc.pf_update(resampler=1, **{k: v for k, v in fi.items() if k != "truth"})


def _check_pairs(got, ora):
    assert got["winner"] == (ora["most_likely_pose_idx"] if ora["resampled"] else ora["most_likely_particle_idx"])
    assert (np.asarray(got["pairs"]).reshape(-1, 2) == ora["winner_pairs"]).all()


def test_config3_full_size_against_the_oracle():
    """1e6 particles x 205 blobs (pose_estimator.cpp:2409-2438 at D = 205): every weight, the stage-wise indices,
    the winner and its correspondences -- not a spot check."""
    N = 1_000_000
    fi = frame_inputs(N, frame=33, n_outliers=200)
    ora = _oracle(fi)
    h = make_handle(fi, N)
    got = run_frame(h, fi)
    res = parity.check_frame(fi, got, ora)
    assert res["index_mismatch"] <= 2
    if res["weight_ties"] == 0 and res["index_mismatch"] == 0:
        _check_pairs(got, ora)
    m, ties, bad = parity.index_report(got, ora, fi["seed"])
    assert not bad, ("index mismatches that are not weight ties", bad[:10])
    h.close()


def test_config4_full_batch_random_trackers_against_the_oracle():
    """The real 4096 x 4096 batch (one launch, one CTA per filter); 16 randomly chosen trackers against the oracle."""
    n_t, N = 4096, 4096
    rng = np.random.default_rng(5)
    pick = sorted(int(t) for t in rng.choice(n_t, 16, replace=False))
    fis = {t: frame_inputs(N, frame=11, tracker=t) for t in pick}
    fi0 = fis[pick[0]]
    h = make_handle(fi0, N, n_trackers=n_t, max_blobs=16)
    assert h.T == n_t
    # every tracker needs a particle set: the picked ones get their own cloud, the others tracker pick[0]'s
    frames = []
    for t in range(n_t):
        fi = fis.get(t, fi0)
        h.set_particles(fi["E"], tracker=t)
        frames.append(fi)
    outs = h.update_batch(frames)
    assert h.launch_info()[:2] == [1, n_t]
    for t in pick:
        got = collect(h, outs[t], tracker=t)
        ora = _oracle(fis[t])
        res = parity.check_frame(fis[t], got, ora)
        if res["weight_ties"] == 0 and res["index_mismatch"] == 0:
            _check_pairs(got, ora)
    # a tracker that was not picked ran the same inputs as pick[0]: bit-identical outputs (independent filters)
    other = next(t for t in range(n_t) if t not in fis)
    a, b = collect(h, outs[pick[0]], tracker=pick[0]), collect(h, outs[other], tracker=other)
    assert (a["w_raw"] == b["w_raw"]).all() and (a["idx"] == b["idx"]).all() and (a["T"] == b["T"]).all()
    h.close()


@pytest.mark.parametrize("W", [2, 8])
def test_config5_full_size_sharded_equals_unsharded_by_checksum(W):
    """2^26 particles: W logical shards against the unsharded filter over three tracked frames, compared by the
    position-keyed device checksums of PoseParticle, the weights and newPoseEstimation (no 7 GB host copy), plus the
    whole result record.  The frame loop is the C one (pfmpe_shard_frame)."""
    N = 1 << 26
    sc = S.Scene(0)
    D = sc.blobs(0).shape[0]
    ref = capi.Handle(N, n_trackers=1, n_leds=5, max_blobs=D + 4)
    ref.configure(sc.markers, sc.K, sc.params)
    ref.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
    sh = sharded.ShardedFilter(N, sharded.Loopback(W), n_leds=5, max_blobs=D + 4)
    assert sh.loop == "c"
    sh.configure(sc.markers, sc.K, sc.params)
    sh.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
    for f in range(2, 5):
        a = ref.track_frame_host(sc.time(f), sc.blobs(f), sc.seed(f))
        b = sh.track_frame_host(sc.time(f), sc.blobs(f), sc.seed(f))
        for k in ("status", "iterations", "most_likely_pose_idx", "max_count", "highest_prob", "weight_sum", "gn_iterations", "flag_fail"):
            assert a[k] == b[k], (f, k, a[k], b[k])
        assert (a["pose"] == b["pose"]).all() and (a["cov"] == b["cov"]).all() and (a["winner_pose"] == b["winner_pose"]).all()
        assert a["most_likely_pose_idx"] >= 0                 # resampled
        assert ref.state_checksum() == sh.state_checksums(), f
    ref.close()
    sh.close()


def test_checksum_detects_a_single_moved_word():
    """The checksum is position keyed: two particles swapped, or one weight changed in its last bit, changes it."""
    N = 4096
    fi = frame_inputs(N, frame=7)
    h = make_handle(fi, N)
    run_frame(h, fi)
    c0 = h.state_checksum()
    E = fi["E"].copy()
    E[[10, 11]] = E[[11, 10]]
    fi2 = dict(fi, E=E)
    run_frame(h, fi2)
    c1 = h.state_checksum()
    assert c0 != c1 and c0[0] != c1[0]
    run_frame(h, fi)
    assert h.state_checksum() == c0                        # and it is reproducible
    h.close()


@pytest.mark.parametrize("N", [100, 10_000, 100_000])
def test_end_to_end_index_agreement(N):
    """north_star: "resampled indices identical except at documented weight ties".  Device indices (FP32 weights) against
    the oracle's (its own FP64 weights) on the same RNG stream; every differing slot must be a weight tie."""
    tot = 0
    for frame in (5, 23):
        fi = frame_inputs(N, frame=frame)
        ora = _oracle(fi)
        h = make_handle(fi, N)
        got = run_frame(h, fi)
        h.close()
        assert got["resampled"] == 1 and ora["resampled"] == 1
        m, ties, bad = parity.index_report(got, ora, fi["seed"])
        assert not bad, ("index mismatches that are not weight ties", N, frame, bad[:10])
        assert m <= max(1, N // 2000), (N, frame, m)      # a handful per 1e5 strata
        tot += m
    print("end-to-end index mismatches N=%d: %d (all weight ties)" % (N, tot))
