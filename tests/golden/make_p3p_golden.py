"""Writes tests/golden/p3p_votes.json: inputs and outputs of the initialisation vote (oracle/p3p_oracle.cpp) for a
few small scenes.  The reference cannot be compiled in this container (Eigen / OpenCV / ROS absent), so these
are ORACLE outputs: they pin the CUDA path and the oracle against regressions, not against the reference."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402
from tests.p3p_util import vote_scene  # noqa: E402

cases = []
for n_out, frame in ((0, 7), (2, 11), (6, 23), (10, 7)):
    sc, img, true_blob = vote_scene(n_out, frame=frame)
    hist, st = orc.p3p_histogram(sc.K, sc.markers, img, 5.0)
    cases.append(dict(n_outliers=n_out, frame=frame, image_points=img.tolist(), true_blob=true_blob,
                      hist=hist.tolist(), stats=list(st)))
json.dump(dict(tol=5.0, cases=cases), open(os.path.join(ROOT, "tests", "golden", "p3p_votes.json"), "w"))
print("wrote", len(cases), "cases")
