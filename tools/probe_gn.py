import sys, os
import numpy as np
sys.path.insert(0, "/root/repo")
from pf_monocular_pose_estimator_b200 import capi, scenes as S
N, T, frames = 4096, 512, 12
h = capi.Handle(N, n_trackers=T, max_blobs=15)
h.configure(S.MARKERS_5, S.K_README, S.DEFAULT_PARAMS)
trackers = np.arange(T); fr = list(range(2, 2 + frames))
blobs, counts = S.blob_tensor_fast(trackers, fr, n_outliers=10)
times = np.array([f * S.DT for f in fr])
seeds = np.array([[S.pf_seed(f, int(t)) for f in fr] for t in trackers], dtype=np.uint32)
for t in trackers:
    ph = S.tracker_phase(int(t)); h.track_init(S.true_pose(0.0, ph), 0.0, S.true_pose(S.DT, ph), S.DT, tracker=int(t))
h.track_upload(blobs, counts, times, seeds)
for f in range(frames):
    h.track_run(f, 1); ms = h.last_launch_ms()
    gn = []; it = []; st = []
    for t in range(T):
        r = h.track_results(f, 1, tracker=t)[0]
        gn.append(r["gn_iterations"]); it.append(r["iterations"]); st.append(r["status"])
    gn = np.array(gn); it = np.array(it)
    print("frame %d: %.3f ms  gn mean %.1f max %d (>=40: %d)  pf iters mean %.2f max %d  status!=1: %d" % (f, ms, gn.mean(), gn.max(), (gn >= 40).sum(), it.mean(), it.max(), (np.array(st) != 1).sum()))
