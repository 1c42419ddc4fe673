"""A few frames of the sharded filter on ONE rank (world = 1) through the C frame loop: ncu target for the stage launches
of pf_frames_kernel<5,1> -- per frame one score-only launch (mode 3), one resample-only launch (mode 4), one winner launch
(mode 5) -- the per-stage counters north_star asks for (FP32 pipe for scoring, HBM GB/s for scan + resample).
    python tools/shard_stage_one.py [particles] [frames]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pf_monocular_pose_estimator_b200 import scenes as S, sharded as SH

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 5
sc = S.Scene(0)
sf = SH.ShardedFilter(N, SH.Loopback(1), n_leds=5, max_blobs=sc.blobs(0).shape[0] + 4)
sf.configure(sc.markers, sc.K, sc.params)
sf.track_init(sc.pose(0), sc.time(0), sc.pose(1), sc.time(1))
for f in range(2, 2 + frames):
    r = sf.track_frame_host(sc.time(f), sc.blobs(f), sc.seed(f))
    print("frame %d status %d iterations %d resampled %d" % (f, r["status"], r["iterations"], r["resampled"]))
sf.close()
