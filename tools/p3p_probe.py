"""GPU probe for row f4: discrepancy statistics of the device P3P / vote against the oracle, and timings."""
import sys
import time

import numpy as np

from oracle import orc
from pf_monocular_pose_estimator_b200 import capi, scenes as S
from tests.p3p_util import random_problems, vote_scene

h = capi.Handle(1000, n_leds=5, max_blobs=32)
h.configure(S.MARKERS_5, S.K_README, S.DEFAULT_PARAMS)

fv, wp, _ = random_problems(20000, seed=9)
rc, sol = h.p3p_solve(fv, wp)
rel = []
nan_mismatch = 0
for i in range(fv.shape[0]):
    r0, s0 = orc.p3p_compute_poses(fv[i], wp[i])
    assert r0 == rc[i]
    a, b = sol[i], s0
    fa, fb = np.isfinite(a), np.isfinite(b)
    nan_mismatch += int((fa != fb).any())
    m = fa & fb
    rel.append(np.max(np.abs(a[m] - b[m]) / (1 + np.abs(b[m]))) if m.any() else 0.0)
rel = np.array(rel)
print("p3p_solve: n=%d nan_pattern_mismatch=%d rel err p50=%.2e p99=%.2e p99.9=%.2e max=%.2e" % (
    len(rel), nan_mismatch, np.percentile(rel, 50), np.percentile(rel, 99), np.percentile(rel, 99.9), rel.max()))

for n_out in (0, 3, 10, 20, 40, 60):
    sc, img, tb = vote_scene(n_out)
    t = time.perf_counter(); ho, so = orc.p3p_histogram(sc.K, sc.markers, img, 5.0); tc = time.perf_counter() - t
    t = time.perf_counter(); hg, sg = h.p3p_histogram(img); tg = time.perf_counter() - t
    hg2, _ = h.p3p_histogram(img)
    d = hg.astype(np.int64) - ho.astype(np.int64)
    hi, shi = orc.p3p_histogram(sc.K, sc.markers, img, 5.0, dup_mode=1)
    lo, slo = orc.p3p_histogram(sc.K, sc.markers, img, 5.0, dup_mode=2)
    below = np.maximum(lo.astype(np.int64) - hg, 0); above = np.maximum(hg.astype(np.int64) - hi, 0)
    print("   bracket: sum(hi-lo)=%d  gpu below lo: %d (cells %d)  above hi: %d (cells %d); ref inside: %s; projected lo/ref/gpu/hi = %d/%d/%d/%d" % (
        int(hi.sum()) - int(lo.sum()), int(below.sum()), int((below > 0).sum()), int(above.sum()), int((above > 0).sum()),
        bool(((lo <= ho) & (ho <= hi)).all()), slo[2], so[2], sg["solutions_projected"], shi[2]))
    print("D=%d votes=%d |diff|=%d (cells %d) stats cpu=%s gpu=%s deterministic=%s cpu %.3fs gpu %.4fs dev %.3f ms" % (
        img.shape[0], int(ho.sum()), int(np.abs(d).sum()), int((d != 0).sum()), so,
        (sg["triples"], sg["p3p_solved"], sg["solutions_projected"], sg["solutions_voted"]),
        bool((hg == hg2).all()), tc, tg, sg["device_ms"]))

if len(sys.argv) > 1:
    for D in (105, 205):
        sc, img, tb = vote_scene(D - 5)
        for rep in range(3):
            t = time.perf_counter(); hg, sg = h.p3p_histogram(img); tg = time.perf_counter() - t
            print("D=%d gpu wall %.4fs dev %.3f ms solves=%d -> %.3e P3P/s; true pairs rank:" % (
                D, tg, sg["device_ms"], sg["p3p_solved"], sg["p3p_solved"] / (sg["device_ms"] * 1e-3)),
                [int((hg[:, j] > hg[tb[j], j]).sum()) for j in range(5)])
