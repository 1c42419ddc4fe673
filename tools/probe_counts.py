"""Scratch probe: offspring-count statistics of the bench sequences (GPU)."""
import sys, numpy as np
sys.path.insert(0, __import__("os").path.dirname(__import__("os").path.dirname(__import__("os").path.abspath(__file__))))
from pf_monocular_pose_estimator_b200 import capi, scenes as S

def run(N, T=1, nout=10, frames=20):
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
    h.track_run(0, frames)
    h.sync()
    for t in range(min(T, 3)):
        c = h.resample_counts(tracker=t)
        w = h.weights(False, tracker=t)
        res = h.track_results(0, frames, tracker=t)
        print("N=%d trk=%d: zero-count %.3f  cnt>4 %.4f  max %d  w==0 %.3f  iters %s gn %s" % (
            N, t, (c == 0).mean(), (c > 4).mean(), c.max(), (w == 0).mean(),
            [r["iterations"] for r in res][-5:], [r["gn_iterations"] for r in res][-5:]))
    if T > 3:
        it = np.array([[r["iterations"] for r in h.track_results(0, frames, tracker=t)] for t in range(T)])
        st = np.array([[r["status"] for r in h.track_results(0, frames, tracker=t)] for t in range(T)])
        print("  all trackers: iterations max %d mean %.3f; status!=1: %d" % (it.max(), it.mean(), (st != 1).sum()))
    h.close()

if __name__ == "__main__":
    run(100_000)
    run(4096, T=256)
