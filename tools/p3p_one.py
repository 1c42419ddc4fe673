"""One initialisation vote (K9) at D detections, for profiling: python tools/p3p_one.py [D] [repeats]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pf_monocular_pose_estimator_b200 import capi
from tests.p3p_util import vote_scene

D = int(sys.argv[1]) if len(sys.argv) > 1 else 205
R = int(sys.argv[2]) if len(sys.argv) > 2 else 3
sc, img, _ = vote_scene(D - 5)
h = capi.Handle(1000, n_leds=5, max_blobs=32)
h.configure(sc.markers, sc.K, sc.params)
ms = []
for _ in range(R):
    hist, st = h.p3p_histogram(img)
    ms.append(st["device_ms"])
print("D=%d vote %.3f ms  solves %d  votes %d" % (D, float(np.median(ms)), st["p3p_solved"], int(hist.sum())))
h.close()
