"""Per-stage timeline of the persistent frame kernel (in-kernel %globaltimer stamps)."""
import sys
import numpy as np
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from pf_monocular_pose_estimator_b200 import capi, scenes as S

NAMES = ["ctx", "blobs", "score(cta0)", "reduce1", "->resample", "resample(cta0)", "reduce2", "result", "tail", "endsync"]


def run(N, T=1, nout=10, frames=40, mode=0):
    h = capi.Handle(N, n_trackers=T, max_blobs=5 + nout, force_mode=mode)
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
    h.debug_timeline(True)
    rows, ms = [], []
    for f in range(frames):
        h.track_run(f, 1)
        ms.append(h.last_launch_ms())
        rows.append(h.debug_timeline(True).astype(np.int64))
    rows = np.array(rows[5:])
    d = np.diff(rows[:, :10], axis=1) / 1e3
    print("N=%d T=%d D=%d grid=%s  launch ms p50=%.4f" % (N, T, 5 + nout, h.launch_info(), np.median(ms[5:])))
    print("  stage us (median): " + "  ".join("%s=%.1f" % (NAMES[i + 1], np.median(d[:, i])) for i in range(9)))
    print("  in-kernel total us: %.1f   GN us: %.1f  predict us: %.1f" % (np.median((rows[:, 9] - rows[:, 0]) / 1e3), np.median((rows[:, 10] - rows[:, 7]) / 1e3), np.median((rows[:, 11] - rows[:, 10]) / 1e3)))
    print("  resample detail us: prefix=%.1f  ktot+kcarry=%.1f  walk=%.1f  sync+list=%.1f" % tuple(
        np.median((rows[:, b] - rows[:, a]) / 1e3) for a, b in ((4, 12), (12, 13), (13, 14), (14, 5))))
    h.close()


if __name__ == "__main__":
    run(100_000)
    run(100_000, mode=1, frames=12)
    run(1_000_000, nout=200, frames=12)
    run(4096, T=1024, frames=12)
    run(1 << 24, frames=10)
