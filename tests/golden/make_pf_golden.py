"""Writes tests/golden/pf_update.json: outputs of one PF update (oracle/pf_oracle.c, pose_estimator.cpp:475-733) for
a few small seeded frames.  The inputs are regenerated from the seeds by tests/util.frame_inputs (a checksum of
them is stored).  The reference cannot be compiled in this container (Eigen / OpenCV / ROS absent), so these are
ORACLE outputs: they pin the oracle, the host-math twin and the CUDA path against regressions."""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import orc  # noqa: E402
from tests.util import frame_inputs  # noqa: E402

CASES = [dict(N=64, frame=5, n_outliers=10, isi=2), dict(N=200, frame=23, n_outliers=10, isi=2),
         dict(N=300, frame=40, n_outliers=40, isi=2), dict(N=128, frame=9, n_outliers=3, isi=1)]


def checksum(fi):
    h = hashlib.sha256()
    for k in ("img", "E", "T_cur", "T_pred", "P"):
        h.update(np.ascontiguousarray(fi[k], dtype=np.float64).tobytes())
    return h.hexdigest()


if __name__ == "__main__":
    out = []
    for c in CASES:
        fi = frame_inputs(c["N"], frame=c["frame"], n_outliers=c["n_outliers"], isi=c["isi"])
        o = orc.pf_update(resampler=0, **{k: v for k, v in fi.items() if k != "truth"})
        out.append(dict(case=c, inputs_sha256=checksum(fi), iterations=int(o["iterations"]), best_iteration=int(o["best_iteration"]),
                        accepted=int(o["accepted"]), resampled=int(o["resampled"]), highest_prob=float(o["highest_prob"]),
                        weight_sum=float(o["weight_sum"]), most_likely_particle_idx=int(o["most_likely_particle_idx"]),
                        most_likely_pose_idx=int(o["most_likely_pose_idx"]), w_raw=[float(x) for x in o["w_raw"]],
                        idx=[int(x) for x in o["idx"]], count=[int(x) for x in o["count"]],
                        winner_pose=np.asarray(o["winner_pose"]).reshape(4, 4).tolist(),
                        T_first=np.asarray(o["T"][:4]).tolist()))
    json.dump(dict(cases=out), open(os.path.join(ROOT, "tests", "golden", "pf_update.json"), "w"))
    print("wrote", len(out), "cases")
