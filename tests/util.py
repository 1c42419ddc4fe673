"""Shared helpers for the parity tests: seeded PF frame inputs built from the synthetic scene."""
import numpy as np

from pf_monocular_pose_estimator_b200 import scenes as S


def f32(x):
    """Round to float32-representable doubles (identical inputs for the FP32 CUDA path and the FP64 oracle)."""
    return np.asarray(x, dtype=np.float64).astype(np.float32).astype(np.float64)


def perturb(T, rng, ang, tr):
    a, b, c = rng.uniform(-ang, ang, 3)
    R = S.rot_z(c) @ S.rot_y(b) @ S.rot_x(a)
    out = T.copy()
    out[:3, :3] = T[:3, :3] @ R
    out[:3, 3] = T[:3, 3] + rng.uniform(-tr, tr, 3)
    return out


def frame_inputs(N, frame=5, tracker=0, n_outliers=10, isi=2, seed=None, spread=(0.004, 0.004), occlude=(),
                 markers=S.MARKERS_5, rng_seed=7):
    """Inputs of one PF update at `frame`: particle cloud around the frame-(f-1) pose."""
    sc = S.Scene(tracker=tracker, n_outliers=n_outliers, markers=markers)
    rng = np.random.default_rng(rng_seed + 31 * frame + 1009 * tracker)
    prev, cur, nxt = sc.pose(frame - 2), sc.pose(frame - 1), sc.pose(frame)
    P = np.linalg.inv(prev) @ cur          # constant-velocity stand-in for predictPose
    T_cur = f32(cur)
    T_pred = f32(cur @ P)
    E = np.stack([perturb(cur, rng, spread[0], spread[1]) for _ in range(N)])
    E = f32(E)
    E[:, 3, :] = [0, 0, 0, 1]
    return dict(params=sc.params, markers=sc.markers, K=sc.K, img=sc.blobs(frame, occlude=occlude), E=E,
                T_cur=T_cur, T_pred=T_pred, P=f32(P), Cm=np.eye(4), isi=isi, dt=S.DT,
                seed=sc.seed(frame) if seed is None else seed, truth=nxt)
