"""Reference-faithful host loop around pfmpe_update (what the reference-side shim does between
setImagePoints() and the write-back, pose_estimator.cpp:223-733 minus detection/ROI): predictPose ->
PF update (device) -> optimiseAndUpdatePose (host FP64) -> jump check.  Used by tests and bench.py."""
import numpy as np

from pf_monocular_pose_estimator_b200 import capi


class HostTracker:
    def __init__(self, h, scene, tracker=0, init_particles=True):
        self.h, self.sc, self.t = h, scene, tracker
        self.prev, self.cur = scene.pose(0), scene.pose(1)
        self.t_prev, self.t_cur = scene.time(0), scene.time(1)
        self.isi = 2
        self.cov = np.zeros((6, 6))
        if init_particles:
            h.set_particles(np.repeat(self.cur[None], h.N, axis=0), tracker=tracker)

    def frame(self, f, img=None, seed=None):
        sc = self.sc
        img = sc.blobs(f) if img is None else img
        seed = sc.seed(f) if seed is None else seed
        t_pred = sc.time(f)
        P, pred = capi.predict_pose(self.prev, self.cur, self.t_prev, self.t_cur, t_pred)      # :234
        out = self.h.update(img, self.cur, pred, P, np.eye(4), self.isi, t_pred - self.t_cur, seed, tracker=self.t)
        flag = float(out["status"])
        if out["resampled"]:
            before = out["winner_pose"]
            pose, cov, it = capi.optimise_pose(sc.K, sc.markers, img, out["correspondences"], before)   # :690
            if self.isi < 2:
                self.isi += 1
            self.prev, self.cur = self.cur, pose                                                  # :2013-2014
            if t_pred - self.t_cur > 0.001 or t_pred < self.t_cur:
                self.t_prev, self.t_cur = self.t_cur, t_pred
            self.cov = cov
            if (np.abs(before[:3, :3] - pose[:3, :3]) >= 0.3).any():
                flag = 1.5
            out["gn_iterations"] = it
        else:
            self.isi = 0
        out["flag_fail"] = flag
        out["pose"] = self.cur.copy()
        out["cov"] = self.cov.copy()
        return out
