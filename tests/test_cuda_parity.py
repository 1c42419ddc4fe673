"""-m gpu: parity of the CUDA path (through the C ABI) with the oracle on identical inputs and an
identical RNG stream.  Tolerances and tie classes: tests/parity.py."""
import numpy as np
import pytest

from oracle import orc
from pf_monocular_pose_estimator_b200 import capi, scenes as S
from tests import parity
from tests.cuda_util import collect, make_handle, run_frame
from tests.util import frame_inputs

pytestmark = pytest.mark.gpu


def _oracle(fi, N):
    return orc.pf_update(resampler=(0 if N <= 1000 else 1), **{k: v for k, v in fi.items() if k != "truth"})


def _check_pairs(got, ora):
    q = got["winner"]
    assert got["pairs"].tolist() == ora["pairs"][q][: 2 * ora["npairs"][q]].reshape(-1, 2).tolist()


# (N, frame, isi, outliers, force_mode, grid_cap): batch mode (one CTA), grid mode with few / many CTAs
CASES = [
    (100, 5, 2, 10, 0, 0),          # config 1: default particle count, one CTA
    (100, 9, 1, 10, 0, 0),          # frame right after initialise(): fac* = 1, no predictionMatrix
    (100, 6, 2, 10, 2, 0),          # same size forced through the cooperative grid path
    (1000, 23, 2, 10, 0, 0),
    (1000, 23, 2, 10, 2, 3),        # 3 CTAs share 4 tiles
    (5000, 31, 2, 10, 1, 0),        # one CTA, 20 score tiles / 5 scan tiles with register carry
    (5000, 31, 2, 10, 2, 0),        # look-back over 5 scan tiles
    (300, 17, 2, 200, 0, 0),        # outlier-heavy (config 3 shape)
    (100000, 40, 2, 10, 0, 0),      # config 2 size, auto geometry
    (5000, 31, 2, 10, 3, 0),        # cluster pair: two 128-thread CTAs share the filter through distributed shared memory
    (100, 5, 2, 10, 3, 0),          # ... with an (almost) empty second CTA
]


@pytest.mark.parametrize("N,frame,isi,nout,mode,cap", CASES)
def test_update_parity(N, frame, isi, nout, mode, cap):
    fi = frame_inputs(N, frame=frame, isi=isi, n_outliers=nout)
    ora = _oracle(fi, N)
    h = make_handle(fi, N, force_mode=mode, grid_cap=cap)
    got = run_frame(h, fi)
    res = parity.check_frame(fi, got, ora)
    assert got["status"] == 1 and got["resampled"] == 1
    _check_pairs(got, ora)
    assert res["index_mismatch"] <= (1 if N >= 100000 else 0)
    np.testing.assert_allclose(got["winner_pose"][:3], got["T"][got["winner"]][:3], atol=0)
    assert abs(got["weight_sum"] - got["w_raw"].sum()) < 1e-6 * N
    h.close()


@pytest.mark.parametrize("mode", [1, 2])
def test_update_multi_iteration(mode):
    # occluded LED + outliers: max score < L*min(5,D) -> all 80 iterations, best iteration regenerated
    N = 600
    fi = frame_inputs(N, frame=11, occlude=(2,))
    ora = _oracle(fi, N)
    assert ora["iterations"] == 80 and ora["best_iteration"] != 79
    h = make_handle(fi, N, force_mode=mode)
    got = run_frame(h, fi)
    parity.check_frame(fi, got, ora)
    _check_pairs(got, ora)
    h.close()


@pytest.mark.parametrize("mode", [1, 2])
def test_update_reject_path(mode):
    N = 400
    fi = frame_inputs(N, frame=7)
    fi["img"] = fi["img"] + 300.0
    ora = _oracle(fi, N)
    h = make_handle(fi, N, force_mode=mode)
    got = run_frame(h, fi)
    assert got["status"] == 4 and got["resampled"] == 0 and got["iterations"] == 80
    parity.check_frame(fi, got, ora)
    with pytest.raises(capi.PfmpeError):
        h.resample_indices()
    h.close()


def test_edge_cases_blobs():
    N = 64
    # no detections at all
    fi = frame_inputs(N, frame=5)
    fi["img"] = np.zeros((0, 2))
    h = make_handle(fi, N, max_blobs=8)
    got = run_frame(h, fi)
    ora = _oracle(fi, N)
    assert got["status"] == 4 and got["iterations"] == ora["iterations"] == 1
    h.close()
    # fewer detections than LEDs (3): min(D, L) rounds, exit threshold L*min(5,3)
    fi = frame_inputs(N, frame=5, n_outliers=0, occlude=(0, 4))
    assert fi["img"].shape[0] == 3
    ora = _oracle(fi, N)
    h = make_handle(fi, N)
    got = run_frame(h, fi)
    parity.check_frame(fi, got, ora)
    h.close()
    # 4-LED marker set (demo_marker_positions.yaml)
    fi = frame_inputs(200, frame=8, markers=S.MARKERS_4)
    ora = _oracle(fi, 200)
    h = make_handle(fi, 200)
    got = run_frame(h, fi)
    parity.check_frame(fi, got, ora)
    h.close()


def test_smallest_filter_and_downgrade():
    fi = frame_inputs(2, frame=5)
    ora = _oracle(fi, 2)
    h = make_handle(fi, 2)
    got = run_frame(h, fi)
    parity.check_frame(fi, got, ora)
    h.close()
    fi = frame_inputs(128, frame=12)
    fi["params"] = dict(fi["params"], bMarkerDowngrade=[0, 1, 0, 0, 1])
    ora = _oracle(fi, 128)
    h = make_handle(fi, 128)
    got = run_frame(h, fi)
    parity.check_frame(fi, got, ora)
    h.close()


@pytest.mark.parametrize("N,mode", [(30000, 1), (30000, 2), (700, 2)])
def test_dominant_particles_fill_tiers(N, mode):
    """Almost every particle misses the gate, so a handful of sources own thousands of output slots each:
    exercises the lane / warp / CTA tiers of the slot fill and the 'not found keeps the previous index' tail."""
    fi = frame_inputs(N, frame=14)
    far = fi["E"][40:].copy()
    far[:, 0, 3] += 5.0                   # projections leave the image: weight 0
    fi["E"][40:] = far
    ora = _oracle(fi, N)
    assert ora["resampled"] == 1 and (ora["w_raw"] > 0).sum() <= 40
    h = make_handle(fi, N, force_mode=mode)
    got = run_frame(h, fi)
    res = parity.check_frame(fi, got, ora)
    assert res["index_mismatch"] == 0
    assert got["count"].max() > (500 if N > 20000 else 10) and got["count"].max() <= 4096
    assert (np.bincount(got["idx"], minlength=N) == got["count"]).all()
    # one source owning (almost) everything: CTA-cooperative tier
    fi["E"][2:] = far[0]
    ora = _oracle(fi, N)
    got = run_frame(h, fi)
    res = parity.check_frame(fi, got, ora)
    assert res["index_mismatch"] == 0 and got["count"].max() > N // 3
    h.close()


def test_huge_slot_ranges_are_filled_by_the_grid():
    """Three survivors of 400k particles own > 65536 output slots each: grid-wide fill after the stage barrier
    (plus the extra barrier before the next frame reads idx)."""
    N = 400000
    fi = frame_inputs(2000, frame=14)
    E = np.repeat(fi["E"][:1], N, axis=0)
    E[:, 0, 3] += 5.0
    E[1000] = fi["E"][3]
    E[1001] = fi["E"][4]
    E[250000] = fi["E"][5]
    fi["E"] = E
    ora = _oracle(fi, N)
    assert ora["resampled"] == 1
    h = make_handle(fi, N)
    got = run_frame(h, fi)
    res = parity.check_frame(fi, got, ora)
    assert res["index_mismatch"] == 0
    assert got["count"].max() > 65536
    assert (np.bincount(got["idx"], minlength=N) == got["count"]).all()
    # the next frame reads its sources through idx[] inside the same tracker state
    out2 = h.update(fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], fi["isi"], fi["dt"], fi["seed"] + 1)
    assert out2["status"] == 1
    h.close()


def test_grid_and_batch_paths_are_bit_identical():
    # exact fixed-point sums + exact thresholds: any decomposition gives the same weights, kend, indices
    N = 3000
    fi = frame_inputs(N, frame=19)
    outs = []
    for mode, cap in [(1, 0), (2, 2), (2, 5), (2, 0), (3, 0)]:
        h = make_handle(fi, N, force_mode=mode, grid_cap=cap)
        outs.append(run_frame(h, fi))
        h.close()
    for o in outs[1:]:
        assert (o["w_raw"] == outs[0]["w_raw"]).all()
        assert (o["T"] == outs[0]["T"]).all()
        assert (o["idx"] == outs[0]["idx"]).all() and (o["count"] == outs[0]["count"]).all()
        assert o["weight_sum"] == outs[0]["weight_sum"] and o["winner"] == outs[0]["winner"]


@pytest.mark.parametrize("N,mode", [(500, 1), (5000, 2)])
def test_second_frame_gathers_through_resample_indices(N, mode):
    """Frame 2 reads its source particles through the previous frame's resampling (fused gather):
    the oracle is fed the resampled set the CUDA path reports."""
    fi1 = frame_inputs(N, frame=20)
    h = make_handle(fi1, N, force_mode=mode)
    got1 = run_frame(h, fi1)
    E2 = h.particles(1)
    np.testing.assert_array_equal(E2, got1["T"][got1["idx"]])
    fi2 = frame_inputs(N, frame=21)
    fi2["E"] = E2
    ora2 = _oracle(fi2, N)
    got2 = run_frame(h, fi2, set_particles=False)
    parity.check_frame(fi2, got2, ora2)
    _check_pairs(got2, ora2)
    h.close()


@pytest.mark.parametrize("N,T,mode,shape", [(512, 5, 0, [1, 5, 256, 0]), (4096, 6, 3, [2, 6, 128, 2])])
def test_update_batch_matches_single_updates(N, T, mode, shape):
    """One launch for all trackers: one CTA per filter, or (force_mode 3) a cluster pair per filter."""
    fis = [frame_inputs(N, frame=10 + t, tracker=t) for t in range(T)]
    hb = make_handle(fis[0], N, n_trackers=T, force_mode=mode)
    for t in range(T):
        hb.set_particles(fis[t]["E"], tracker=t)
    outs = hb.update_batch([dict(img=f["img"], T_cur=f["T_cur"], T_pred=f["T_pred"], P=f["P"], Cm=f["Cm"], isi=f["isi"],
                                 dt=f["dt"], seed=f["seed"]) for f in fis])
    assert hb.launch_info() == shape
    for t in range(T):
        got = collect(hb, outs[t], tracker=t)
        ora = _oracle(fis[t], N)
        parity.check_frame(fis[t], got, ora)
        hs = make_handle(fis[t], N)
        single = run_frame(hs, fis[t])
        assert (single["w_raw"] == got["w_raw"]).all() and (single["idx"] == got["idx"]).all()
        hs.close()
    hb.close()


def test_call_sequence_errors():
    fi = frame_inputs(64, frame=5)
    h = capi.Handle(64, max_blobs=4)
    with pytest.raises(capi.PfmpeError) as e:
        h.update(fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], 2, 0.02, 1)
    assert e.value.code == capi.ERR_STATE
    h.configure(fi["markers"], fi["K"], fi["params"])
    with pytest.raises(capi.PfmpeError) as e:
        h.update(fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], 2, 0.02, 1)
    assert e.value.code == capi.ERR_STATE          # set_particles missing
    h.set_particles(fi["E"])
    with pytest.raises(capi.PfmpeError) as e:
        h.update(fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], 2, 0.02, 1)
    assert e.value.code == capi.ERR_INVALID        # 15 detections > max_blobs 4
    h.close()


def test_full_size_properties():
    """BASELINE config-3 size (1e6 particles, 205 blobs): size-independent invariants."""
    N = 1_000_000
    fi = frame_inputs(N, frame=33, n_outliers=200)
    h = make_handle(fi, N)
    got = run_frame(h, fi)
    assert got["status"] == 1 and got["resampled"] == 1
    idx, cnt, w = got["idx"], got["count"], got["w_raw"]
    assert cnt.sum() == N                                 # every stratum found a source
    assert (np.diff(idx) >= 0).all()                      # stratified indices are sorted
    assert (np.bincount(idx, minlength=N) == cnt).all()   # counts are the histogram of indices
    assert (cnt[w <= 0] == 0).all() or (w < 0).any()
    assert got["winner"] == int(np.argmax(cnt)) and got["max_count"] == cnt.max()
    assert abs(got["weight_sum"] - w.astype(np.float64).sum()) < 1e-3
    assert w.max() <= 30.0 + 1e-4 and abs(got["highest_prob"] - w.max()) < 1e-6
    # expected offspring ~ N * w / S within one stratum
    exp = N * w / got["weight_sum"]
    assert np.abs(cnt - exp).max() < 2.0
    # (every weight of this configuration is compared with the oracle in tests/test_cuda_fullsize.py)
    h.close()


@pytest.mark.parametrize("N,mode", [(100, 0), (30000, 0), (5000, 1)])
def test_weighted_mean_pose_and_covariance(N, mode):
    """K7 (north-star extra; no reference counterpart): weighted mean pose + 6x6 covariance of
    the kept iteration's particles, against the CPU restatement fed with the CUDA path's own T and weights."""
    fi = frame_inputs(N, frame=11)
    h = make_handle(fi, N, flags=capi.FLAG_MEAN_COV, force_mode=mode)
    got = run_frame(h, fi)
    assert got["resampled"] == 1
    w = got["w_raw"] / got["weight_sum"]
    mean, cov = orc.weighted_mean_cov(got["T"], w)
    np.testing.assert_allclose(got["mean_pose"], mean, rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(got["mean_cov"], cov, rtol=1e-7, atol=1e-16)
    R = got["mean_pose"][:3, :3]
    assert abs(np.linalg.det(R) - 1) < 1e-12 and np.abs(R @ R.T - np.eye(3)).max() < 1e-12
    assert np.abs(got["mean_pose"][:3, 3] - fi["truth"][:3, 3]).max() < 0.1
    assert np.all(np.linalg.eigvalsh(got["mean_cov"]) > -1e-15)
    # without the flag the fields stay zero
    h2 = make_handle(fi, N)
    g2 = run_frame(h2, fi)
    assert not g2["mean_cov"].any()
    h.close(); h2.close()


def test_weighted_mean_cov_batched_trackers():
    N, T = 2048, 3
    fis = [frame_inputs(N, frame=6 + t, tracker=t) for t in range(T)]
    h = make_handle(fis[0], N, n_trackers=T, flags=capi.FLAG_MEAN_COV)
    for t in range(T):
        h.set_particles(fis[t]["E"], tracker=t)
    outs = h.update_batch([dict(img=f["img"], T_cur=f["T_cur"], T_pred=f["T_pred"], P=f["P"], Cm=f["Cm"], isi=f["isi"],
                                dt=f["dt"], seed=f["seed"]) for f in fis])
    for t in range(T):
        Tt, w = h.particles(0, tracker=t), h.weights(False, tracker=t) / outs[t]["weight_sum"]
        mean, cov = orc.weighted_mean_cov(Tt, w)
        np.testing.assert_allclose(outs[t]["mean_pose"], mean, rtol=1e-9, atol=1e-12)
        np.testing.assert_allclose(outs[t]["mean_cov"], cov, rtol=1e-7, atol=1e-16)
    h.close()


@pytest.mark.parametrize("N,mode", [(100, 0), (20000, 0), (3000, 1)])
def test_roi_prediction_matches_oracle(N, mode):
    """K8 (row f1): bounding corners of all resampled particles x LEDs + predicted pose, against
    predictMarkerPositionsInImage :1036-1052 + determineROI's extreme-value loop (led_detector.cpp:221-244)."""
    fi = frame_inputs(N, frame=12)
    h = make_handle(fi, N, force_mode=mode)
    h.set_particles(fi["E"])
    Cm = np.eye(4); Cm[0, 3] = 0.01
    args = (Cm, fi["P"], fi["T_pred"])
    # before any update: newPoseEstimation is what initialise() left behind
    np.testing.assert_allclose(h.predict_roi(*args), orc.predict_roi(fi["E"], Cm, fi["P"], fi["T_pred"], fi["markers"], fi["K"]),
                               rtol=1e-5, atol=2e-3)
    got = run_frame(h, fi, set_particles=False)
    assert got["resampled"] == 1
    roi = h.predict_roi(*args)                      # through idx[]: the resampled set is never materialised
    ref = orc.predict_roi(h.particles(1), Cm, fi["P"], fi["T_pred"], fi["markers"], fi["K"])
    np.testing.assert_allclose(roi, ref, rtol=1e-5, atol=2e-3)
    assert roi[0] < roi[2] and roi[1] < roi[3]
    # the reference's start values: x_max / y_max never drop below 0, x_min / y_min follow the data
    E = fi["E"].copy(); E[:, 0, 3] -= 5.0; E[:, 1, 3] -= 5.0
    h.set_particles(E)
    pred = fi["T_pred"].copy(); pred[0, 3] -= 5.0; pred[1, 3] -= 5.0
    roi = h.predict_roi(np.eye(4), np.eye(4), pred)
    ref = orc.predict_roi(E, np.eye(4), np.eye(4), pred, fi["markers"], fi["K"])
    assert ref[2] == 0.0 and ref[3] == 0.0 and roi[2] == 0.0 and roi[3] == 0.0
    np.testing.assert_allclose(roi, ref, rtol=1e-5, atol=2e-3)
    h.close()


@pytest.mark.parametrize("N,mode", [(100, 0), (20000, 0), (2000, 1)])
def test_plumb_bob_projection_option(N, mode):
    """Row f3 (PFMPE_FLAG_DISTORT): particles are projected through K and the plumb_bob model of
    LEDDetector::distortPoints (led_detector.cpp:371-414) and scored against the raw blob centres."""
    D5 = np.array([-0.28, 0.09, 1.2e-3, -7e-4, 0.01])
    fi = frame_inputs(N, frame=13)
    fi["img"] = orc.distort_points(fi["K"], D5, fi["img"])
    fi["D5"] = D5
    ora = orc.pf_update(resampler=(0 if N <= 1000 else 1), dist=D5, **{k: v for k, v in fi.items() if k not in ("truth", "D5")})
    assert ora["resampled"] == 1
    h = make_handle(fi, N, flags=capi.FLAG_DISTORT, force_mode=mode)
    got = run_frame(h, fi)
    res = parity.check_frame(fi, got, ora)
    assert res["index_mismatch"] == 0
    _check_pairs(got, ora)
    # the same raw detections through the undistorted model give different scores: the option is doing something
    h0 = make_handle(fi, N, force_mode=mode)
    g0 = run_frame(h0, fi)
    assert np.abs(g0["w_raw"] - got["w_raw"]).max() > 1e-3
    # the device-resident tracking modes need undistorted points for Gauss-Newton: refused, not silently wrong
    with pytest.raises(capi.PfmpeError):
        h.track_init(fi["T_cur"], 0.0, fi["T_pred"], 0.02)
        sc = S.Scene(0)
        h.track_upload(np.zeros((1, 1, h.max_blobs, 2)), np.zeros((1, 1), dtype=np.int32), np.array([0.04]), np.ones((1, 1), dtype=np.uint32))
        h.track_run(0, 1)
    h.close(); h0.close()
