"""One initialisation vote on the GPU (for ncu): python tools/p3p_one.py D [reps]"""
import sys

from pf_monocular_pose_estimator_b200 import capi, scenes as S
from tests.p3p_util import vote_scene

D = int(sys.argv[1]) if len(sys.argv) > 1 else 105
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
h = capi.Handle(1000, n_leds=5, max_blobs=32)
h.configure(S.MARKERS_5, S.K_README, S.DEFAULT_PARAMS)
sc, img, tb = vote_scene(D - 5)
for _ in range(reps):
    hist, st = h.p3p_histogram(img)
    print("D=%d dev %.3f ms solves=%d -> %.3e P3P/s votes=%d" % (D, st["device_ms"], st["p3p_solved"],
                                                             st["p3p_solved"] / (st["device_ms"] * 1e-3), int(hist.sum())))
