"""Helpers for the -m gpu tests: run one PF update through the C ABI and collect the same
dictionary of outputs that tests/parity.py checks."""
import numpy as np

from pf_monocular_pose_estimator_b200 import capi


def make_handle(fi, N, n_trackers=1, max_blobs=None, **kw):
    D = np.asarray(fi["img"]).reshape(-1, 2).shape[0]
    h = capi.Handle(N, n_trackers=n_trackers, n_leds=fi["markers"].shape[0], max_blobs=max_blobs or max(D, 1), **kw)
    h.configure(fi["markers"], fi["K"], fi["params"], fi.get("D5"))
    return h


def run_frame(h, fi, tracker=0, set_particles=True):
    if set_particles:
        h.set_particles(fi["E"], tracker=tracker)
    out = h.update(fi["img"], fi["T_cur"], fi["T_pred"], fi["P"], fi["Cm"], fi["isi"], fi["dt"], fi["seed"], tracker=tracker)
    return collect(h, out, tracker)


def collect(h, out, tracker=0):
    got = dict(out)
    got["accepted"] = 1 if out["status"] == 1 else 0
    got["T"] = h.particles(0, tracker=tracker)
    got["w_raw"] = h.weights(False, tracker=tracker)
    got["pairs"] = out["correspondences"]
    if out["resampled"]:
        got["idx"] = h.resample_indices(tracker=tracker)
        got["count"] = h.resample_counts(tracker=tracker)
        got["winner"] = out["most_likely_pose_idx"]
    else:
        got["winner"] = out["most_likely_particle_idx"]
    return got
