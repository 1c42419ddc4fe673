# K9 vote kernel: parity tests, then the D = 205 vote against the resident-CTA setting (PFMPE_P3P_OCC)
export PYTHONPATH=.
mkdir -p gpurun_out
true
for occ in 5 6 8; do
PFMPE_P3P_OCC=$occ python - <<PY
import numpy as np, time
from pf_monocular_pose_estimator_b200 import capi, scenes as S
from tests.p3p_util import vote_scene
for nd in (205, 45):
    sc, img, _ = vote_scene(nd - 5)
    h = capi.Handle(1000, n_leds=5, max_blobs=32)
    h.configure(sc.markers, sc.K, sc.params)
    ms = []
    for i in range(5):
        hist, st = h.p3p_histogram(img)
        ms.append(st["device_ms"])
    print("occ $occ D=%d: %.3f ms (min %.3f)  votes %d  solved %d" % (nd, np.median(ms[1:]), min(ms), int(hist.sum()), st["p3p_solved"]))
    h.close()
PY
done
