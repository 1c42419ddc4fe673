"""Independent NumPy/Python twin of the CPU ORACLE (small cases only).

TEST INFRASTRUCTURE ONLY -- cross-checks oracle/pf_oracle.c.  Written separately from the C
restatement (different data structures: row-indexed numpy 4x4, the "sorted per-LED minima"
form of the score from SURVEY A.4 instead of the literal greedy matrix loop) so that an
error in one restatement shows up as a disagreement.  Cites
/root/reference/pf_mpe_lib/src/pose_estimator.cpp.  Parity status: see pf_oracle.h.
"""
import math

import numpy as np

M = 2147483647
A = 16807
R = 2147483646.0


class MinStd:
    """std::minstd_rand0 + uniform_real_distribution<double> (:476-477, :510-523)."""

    def __init__(self, seed):
        x = (seed & 0xFFFFFFFF) % M
        self.x = x if x else 1
        self.steps = 0

    def next(self):
        self.x = (self.x * A) % M
        self.steps += 1
        return self.x

    def canonical(self):
        g1 = self.next() - 1
        g2 = self.next() - 1
        s = np.float64(g1) * np.float64(1.0)
        s = s + np.float64(g2) * np.float64(R)
        c = s / (np.float64(R) * np.float64(R))
        if c >= 1.0:
            c = np.nextafter(np.float64(1.0), np.float64(0.0))
        return float(c)

    def uniform(self, a, b):
        return self.canonical() * (b - a) + a


def project2d(K, T, X3):
    """:1017-1034 -- p = K (R X + t); (u,v) = p[:2]/p[2]."""
    p = K @ (T[:3, :3] @ X3 + T[:3, 3])
    return p[:2] / p[2]


def score(params, L, img, proj):
    """:2385-2445 in the equivalent parallel form: per-LED nearest blob, then
    LEDs in ascending (min, j) order, at most min(D, L), stop at first beyond tol_PF."""
    D = img.shape[0]
    if D == 0:
        return 0.0, np.zeros((0, 2), dtype=np.uint32)
    tol = params["back_projection_pixel_tolerance"]
    tol_pf = params["back_projection_pixel_tolerance_PF"]
    dg = params.get("bMarkerDowngrade", [0] * 5)
    d2 = ((img[:, None, :] - proj[None, :, :]) ** 2)
    d2 = d2[:, :, 0] + d2[:, :, 1]           # (D, L)
    arg = np.argmin(d2, axis=0)              # first minimum along blobs
    mn = d2[arg, np.arange(L)]
    order = sorted(range(L), key=lambda j: (mn[j], j))
    P = 0.0
    pairs = []
    used = []
    occ = 1
    for j in order[: min(D, L)]:
        d = math.sqrt(mn[j])
        if d <= tol_pf:
            P += L + ((tol - d) / tol) ** 2
            pairs.append((j + 1, int(arg[j]) + 1))
            if int(arg[j]) in used:
                P -= 3 * occ
                occ += 1
            used.append(int(arg[j]))
            if j < 5 and dg[j]:
                P -= 2
        else:
            break
    return P, np.array(pairs, dtype=np.uint32).reshape(-1, 2)


def _rx(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[1, 0, 0, 0], [0, c, -s, 0], [0, s, c, 0], [0, 0, 0, 1.0]])


def _ry(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, 0, s, 0], [0, 1, 0, 0], [-s, 0, c, 0], [0, 0, 0, 1.0]])


def _rz(a):
    c, s = math.cos(a), math.sin(a)
    return np.array([[c, -s, 0, 0], [s, c, 0, 0], [0, 0, 1, 0], [0, 0, 0, 1.0]])


def pf_update(params, markers, K, img, E, T_cur, T_pred, P, Cm, isi, dt, seed, max_iter=80):
    """:475-733.  E: (N,4,4).  Literal O(N^2) resampling."""
    markers = np.asarray(markers, dtype=np.float64)
    img = np.asarray(img, dtype=np.float64).reshape(-1, 2)
    N, L, D = E.shape[0], markers.shape[0], img.shape[0]
    g = MinStd(seed)
    if isi == 1:
        fT = fR = 1.0
    else:
        fT = min(max(0.2, abs(P[0, 3]) / dt), 1.0) / 4
        fR = 0.2
    ta, tb = params["minTransitionNoise"] * fT, params["maxTransitionNoise"] * fT
    aa, ab = params["minAngularNoise"] * fR, params["maxAngularNoise"] * fR
    highest = 0.0
    likely = None
    it = 0
    arg_best = -1
    while True:
        T = np.zeros((N, 4, 4))
        w = np.zeros(N)
        corr = []
        s = 1 + 0.025 * math.floor(it // 10)
        for n in range(N):
            if n == 0:
                T[n] = T_cur
            elif n == 1:
                T[n] = T_pred
            else:
                if isi > 1 and it % 10 != 0:
                    B = (Cm @ E[n]) @ P
                elif isi > 1:
                    B = Cm @ E[n]
                else:
                    B = E[n].copy()
                a = g.uniform(aa, ab) * s
                b = g.uniform(aa, ab) * s
                c = g.uniform(aa, ab) * s
                Tn = ((B @ _rz(c)) @ _ry(b)) @ _rx(a)
                Tn[0, 3] = B[0, 3] + g.uniform(ta, tb) * s
                Tn[1, 3] = B[1, 3] + g.uniform(ta, tb) * s
                Tn[2, 3] = B[2, 3] + g.uniform(ta, tb) * s
                T[n] = Tn
            proj = np.array([project2d(K, T[n], markers[j]) for j in range(L)])
            w[n], pr = score(params, L, img, proj)
            corr.append(pr)
        it += 1
        m = w.max()
        if m > highest:
            highest = m
            arg_best = int(np.argmax(w))
            likely = (T, w, corr, it - 1)
        if not (it < max_iter and m < L * min(5, D)):
            break
    best_it = -1
    if likely is not None:
        T, w, corr, best_it = likely
    w_raw = w.copy()
    S = 0.0
    for v in w:
        S += v
    if S != 0:
        w = w / S
    out = dict(T=T, w_raw=w_raw, w=w, corr=corr, iterations=it, best_iteration=best_it, highest_prob=highest,
               weight_sum=S, most_likely_particle_idx=arg_best, resampled=0, idx=None, count=None,
               most_likely_pose_idx=-1, rng_steps=None)
    if S != 0 and highest > L * min(3, D):
        out["accepted"] = 1
        out["flag_fail"] = 1.0
        if isi > 0:
            idx = np.full(N, -1, dtype=np.int64)
            count = np.zeros(N, dtype=np.int64)
            pi = -1
            for k in range(N):
                u = (k + g.uniform(0.0, 1.0)) / N
                c = 0.0
                for i in range(N):
                    c += w[i]
                    if c >= u:
                        pi = i
                        count[i] += 1
                        break
                idx[k] = pi
            out["idx"], out["count"], out["resampled"] = idx, count, 1
            out["most_likely_pose_idx"] = int(np.argmax(count))
    else:
        out["accepted"] = 0
        out["flag_fail"] = 4.0
        out["most_likely_particle_idx"] = int(np.argmax(w))
    out["rng_steps"] = g.steps
    return out
