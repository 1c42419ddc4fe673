"""The faithful CPU baseline of BASELINE.md section 4: the oracle port of the reference PF with the reference's LITERAL
O(N^2) resampling scan (pose_estimator.cpp:668-682), one thread (the reference is single-threaded), double precision.
  config 1 (100 particles): 1000 tracked frames;  config 2 (1e5 particles): a 10-frame sample.
Prints one JSON line per config: p50 ms per frame, PME/s, CPU model, and beside it the O(N) resampler on the same frames.
    python tools/cpu_faithful.py [frames_cfg1] [frames_cfg2]"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import orc
from pf_monocular_pose_estimator_b200 import scenes as S


def cpu_model():
    try:
        for l in open("/proc/cpuinfo"):
            if l.startswith("model name"):
                return l.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def run(n_p, frames, resampler):
    sc = S.Scene(0)
    trk = orc.OracleTracker(n_p, sc.markers, sc.K, sc.params, sc.pose(0), sc.time(0), sc.pose(1), sc.time(1), resampler=resampler)
    lat, pme, idx_sum = [], 0, 0
    for f in range(2, 2 + frames):
        a = time.perf_counter()
        trk.frame(sc.time(f), sc.blobs(f), sc.seed(f))
        lat.append(time.perf_counter() - a)
        pme += n_p * 5 * trk.pf.iterations
        idx_sum += int(trk.pf.most_likely_pose_idx)
    pose = np.array(trk.current_pose)
    trk.close()
    return dict(p50_ms=1e3 * float(np.median(lat)), pme_per_s=pme / sum(lat), seconds=sum(lat), idx_sum=idx_sum, pose=pose)


def main():
    f1 = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    f2 = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    for name, n_p, frames in (("cfg1", 100, f1), ("cfg2", 100_000, f2)):
        lit = run(n_p, frames, 0)
        lin = run(n_p, frames, 1)
        same = lit["idx_sum"] == lin["idx_sum"] and bool((lit["pose"] == lin["pose"]).all())
        print(json.dumps({"config": name, "particles": n_p, "frames": frames, "threads": 1, "cpu": cpu_model(), "nproc": os.cpu_count(),
                          "literal_O(N^2)": {"p50_ms": round(lit["p50_ms"], 4), "PME_per_s": round(lit["pme_per_s"]), "seconds": round(lit["seconds"], 2)},
                          "O(N)": {"p50_ms": round(lin["p50_ms"], 4), "PME_per_s": round(lin["pme_per_s"]), "seconds": round(lin["seconds"], 2)},
                          "same_winners_and_pose": same}), flush=True)


if __name__ == "__main__":
    main()
