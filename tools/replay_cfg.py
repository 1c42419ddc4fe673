"""PF updates alone (replay mode, the bench's `value` path) for one configuration: device ms per frame.
    python tools/replay_cfg.py N [T] [outliers] [frames] [flush]
An untimed tracked run logs the per-frame constants, then the frames are replayed one launch each (what bench.py
times).  ncu target: the replay launches are the last `frames` pf_frames_kernel launches of the process."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pf_monocular_pose_estimator_b200 import capi, scenes as S

N = int(sys.argv[1]); T = int(sys.argv[2]) if len(sys.argv) > 2 else 1
nout = int(sys.argv[3]) if len(sys.argv) > 3 else 10
frames = int(sys.argv[4]) if len(sys.argv) > 4 else 8
flush = int(sys.argv[5]) if len(sys.argv) > 5 else 0
h = capi.Handle(N, n_trackers=T, max_blobs=5 + nout)
h.configure(S.MARKERS_5, S.K_README, S.DEFAULT_PARAMS)
trackers = np.arange(T)
fr = list(range(2, 2 + frames))
blobs, counts = S.blob_tensor_fast(trackers, fr, n_outliers=nout)
times = np.array([f * S.DT for f in fr])
seeds = np.array([[S.pf_seed(f, int(t)) for f in fr] for t in trackers], dtype=np.uint32)


def init():
    for t in trackers:
        ph = S.tracker_phase(int(t))
        h.track_init(S.true_pose(0.0, ph), 0.0, S.true_pose(S.DT, ph), S.DT, tracker=int(t))


init()
h.track_upload(blobs, counts, times, seeds)
for f in range(frames):
    h.track_run(f, 1)
h.sync()
init()
buf = None
if flush:
    import torch
    buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ms = []
for f in range(frames):
    if buf is not None:
        buf.zero_()
        torch.cuda.synchronize()
    h.track_replay(f, 1)
    ms.append(h.last_launch_ms())
it = np.mean([r["iterations"] for r in h.track_results(0, frames)])
med = float(np.median(ms[2:]))
print("replay N=%d T=%d D=%d grid=%s ms/frame: %s | p50 %.4f ms = %.3e PME/s, %.3f of 6546.9 GB/s on 136 B/particle (mean iterations %.2f)" % (
    N, T, 5 + nout, h.launch_info(), " ".join("%.4f" % m for m in ms), med, N * T * 5 * it / (med * 1e-3),
    N * T * (104 * it + 32) / (med * 1e-3) / 1e9 / 6546.9, it))
h.close()
