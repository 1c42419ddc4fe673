"""Parity checks shared by the host-math (CPU) and CUDA (GPU) tests: compare one frame's outputs
of the FP32 device algorithm with the FP64 oracle on identical inputs and an identical RNG stream.

Tolerances (BASELINE.json north_star): poses and weights within 1e-5 relative in FP32; resampled
indices identical except at documented ties.  Documented tie classes:
  * gate tie   : an LED's nearest-blob distance within GATE_EPS px of tol_PF (the accept gate at
                 pose_estimator.cpp:2414 flips, the score jumps by ~L);
  * argmin tie : two blobs equidistant from a projected LED within ARG_EPS px;
  * stratum tie: u_k within STRATUM_EPS (relative) of a prefix boundary.
"""
import numpy as np

from oracle import orc
from pf_monocular_pose_estimator_b200 import scenes as S

POSE_ATOL = 2e-6         # rotation entries are O(1), translations O(1 m): 1e-5 relative leaves 5x margin
W_RTOL = 1e-5
GATE_EPS = 2e-3
ARG_EPS = 2e-3


def _near_tie_particles(fi, T_oracle):
    """Particles whose oracle score sits within GATE_EPS/ARG_EPS of a discontinuity."""
    img = np.asarray(fi["img"]).reshape(-1, 2)
    tol_pf = fi["params"]["back_projection_pixel_tolerance_PF"]
    N = T_oracle.shape[0]
    flagged = np.zeros(N, dtype=bool)
    if img.shape[0] == 0:
        return flagged
    for n in range(N):
        uv = S.project(fi["K"], T_oracle[n], fi["markers"])
        if fi.get("D5") is not None:
            uv = orc.distort_points(fi["K"], fi["D5"], uv)
        d = np.sqrt(((img[:, None, :] - uv[None, :, :]) ** 2).sum(-1))   # (D, L)
        ds = np.sort(d, axis=0)
        if (np.abs(ds[0] - tol_pf) < GATE_EPS).any():
            flagged[n] = True
        if ds.shape[0] > 1 and ((ds[1] - ds[0] < ARG_EPS) & (ds[0] <= tol_pf + GATE_EPS)).any():
            flagged[n] = True
    return flagged


def check_frame(fi, got, ora, stage_resample=True):
    """got: outputs of the device algorithm (hostmath.frame or the CUDA path); ora: orc.pf_update dict."""
    N = ora["T"].shape[0]
    assert got["iterations"] == ora["iterations"], (got["iterations"], ora["iterations"])
    assert got["best_iteration"] == ora["best_iteration"]
    # propagated poses
    np.testing.assert_allclose(got["T"][:, :3, :], ora["T"][:, :3, :], rtol=0, atol=POSE_ATOL)
    # raw weights: 1e-5 relative except documented ties
    w, wo = got["w_raw"], ora["w_raw"]
    bad = np.abs(w - wo) > W_RTOL * np.maximum(np.abs(wo), 1.0)
    if bad.any():
        ties = _near_tie_particles(fi, ora["T"])
        assert not (bad & ~ties).any(), ("weight mismatch outside tie classes", np.nonzero(bad & ~ties)[0][:10],
                                        w[bad & ~ties][:5], wo[bad & ~ties][:5])
    n_tie = int(bad.sum())
    assert n_tie <= max(2, N // 200), "too many gate/argmin ties: %d of %d" % (n_tie, N)
    assert got["accepted"] == ora["accepted"] and got["resampled"] == ora["resampled"]
    # highest weight / its particle: waived only if one of the two argmax candidates is itself a flagged tie particle
    a_g, a_o = got["most_likely_particle_idx"], ora["most_likely_particle_idx"]
    top_is_tie = bool(n_tie) and a_g >= 0 and a_o >= 0 and bool(bad[a_g] or bad[a_o])
    if not top_is_tie:
        assert abs(got["highest_prob"] - ora["highest_prob"]) <= W_RTOL * max(1.0, abs(ora["highest_prob"]))
        if a_g != a_o and a_g >= 0 and a_o >= 0:
            # two particles whose oracle weights agree within the FP32 tolerance may swap places
            assert abs(wo[a_g] - wo[a_o]) <= 2 * W_RTOL * max(1.0, abs(wo[a_o])), (a_g, a_o, wo[a_g], wo[a_o])
        else:
            assert a_g == a_o
    out = dict(weight_ties=n_tie, index_mismatch=0)
    if ora["resampled"] and stage_resample:
        # stage-wise: feed the oracle's literal double resampler the device's own FP32 weights and
        # the reference RNG stream position -> indices must be identical (except stratum ties)
        w32 = got["w_raw"].astype(np.float64)
        s = 0.0
        for v in w32:
            s += v
        r = orc.resample_uniforms(fi["seed"], N, 12 * (N - 2) * ora["iterations"])
        idx_o, cnt_o = orc.resample(w32 / s, r, literal=(N <= 3000))
        mism = np.nonzero(idx_o != got["idx"])[0]
        out["index_mismatch"] = len(mism)
        if len(mism):
            c = np.cumsum(w32 / s)
            u = (np.arange(N) + r) / N
            for k in mism:
                lo, hi = sorted((idx_o[k], got["idx"][k]))
                assert np.abs(c[lo:hi + 1] - u[k]).min() < 1e-9 * max(1.0, N / 1e4), ("index mismatch not at a stratum tie", k)
        else:
            assert (cnt_o == got["count"]).all()
            q = int(np.argmax(cnt_o))
            assert got["winner"] == q
    return out


def index_report(got, ora, seed, slack=1e-12):
    """End-to-end index agreement (SURVEY 7.3 mitigation (b)): the device's resampled indices against the oracle's, each
    computed from its OWN weights (FP32 on the device, FP64 in the oracle), on the same RNG stream.  A slot k whose two
    indices differ, a < b, is a *weight tie* if every cumulative-weight boundary c_i, a <= i < b, that the two sides
    place on different sides of u_k lies within the distance the two weight vectors themselves differ there:
    |c_i - u_k| <= |c_i(device) - c_i(oracle)| + slack.  Returns (mismatches, ties, unexplained slots)."""
    N = ora["T"].shape[0]
    w_g = got["w_raw"].astype(np.float64)
    w_o = ora["w_raw"].astype(np.float64)
    c_g = np.cumsum(w_g / w_g.sum())
    c_o = np.cumsum(w_o / w_o.sum())
    r = orc.resample_uniforms(seed, N, 12 * (N - 2) * ora["iterations"])
    u = (np.arange(N) + r) / N
    mism = np.nonzero(got["idx"] != ora["idx"])[0]
    unexplained = []
    for k in mism:
        a, b = sorted((int(got["idx"][k]), int(ora["idx"][k])))
        i = np.arange(a, b)
        d = np.minimum(np.abs(c_g[i] - u[k]), np.abs(c_o[i] - u[k]))
        if not (d <= np.abs(c_g[i] - c_o[i]) + slack).all():
            unexplained.append(int(k))
    return len(mism), len(mism) - len(unexplained), unexplained

