"""Run a few device-resident frames of one configuration (profiling target for ncu)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pf_monocular_pose_estimator_b200 import capi, scenes as S

N = int(sys.argv[1]); T = int(sys.argv[2]) if len(sys.argv) > 2 else 1
nout = int(sys.argv[3]) if len(sys.argv) > 3 else 10
frames = int(sys.argv[4]) if len(sys.argv) > 4 else 8
h = capi.Handle(N, n_trackers=T, max_blobs=5 + nout)
h.configure(S.MARKERS_5, S.K_README, S.DEFAULT_PARAMS)
trackers = np.arange(T)
fr = list(range(2, 2 + frames))
blobs, counts = S.blob_tensor_fast(trackers, fr, n_outliers=nout)
times = np.array([f * S.DT for f in fr])
seeds = np.array([[S.pf_seed(f, int(t)) for f in fr] for t in trackers], dtype=np.uint32)
for t in trackers:
    ph = S.tracker_phase(int(t))
    h.track_init(S.true_pose(0.0, ph), 0.0, S.true_pose(S.DT, ph), S.DT, tracker=int(t))
h.track_upload(blobs, counts, times, seeds)
ms = []
for f in range(frames):
    h.track_run(f, 1)
    ms.append(h.last_launch_ms())
print("N=%d T=%d D=%d launch ms: %s" % (N, T, 5 + nout, " ".join("%.4f" % m for m in ms)))
h.close()
