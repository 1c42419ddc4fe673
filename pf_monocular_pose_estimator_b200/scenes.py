"""Synthetic scenes for the PF hot path.

Fixture data comes from the reference README: the 5-LED marker set
(/root/reference/README.md:102-117), the 752x480 camera_info (README.md:137-141) and the
PF launch parameters (README.md:338,356-361).  The trajectory / blob generator is ours
(seeded numpy), used identically by the CUDA path, the oracle and the CPU baseline.
Host-side numpy only; nothing here touches the GPU.
"""
import numpy as np

# README.md:102-117 (metres, object frame)
MARKERS_5 = np.array([
    [0.0, 0.0, 0.0],
    [0.004, -0.188, 0.039],
    [0.228, -0.140, 0.0],
    [0.264, 0.124, 0.0],
    [0.076, 0.128, 0.005],
], dtype=np.float64)

# pf_mpe/marker_positions/demo_marker_positions.yaml:3-15 (4 LEDs)
MARKERS_4 = np.array([
    [0.0714197, 0.0800214, 0.0622611],
    [0.0400755, -0.0912328, 0.0317064],
    [-0.0647293, -0.0879977, 0.0830852],
    [-0.0558663, -0.0165446, 0.053473],
], dtype=np.float64)

IMAGE_W, IMAGE_H = 752, 480
# README.md:137-141, row-major 3x3 as in sensor_msgs/CameraInfo.K
K_README = np.array([
    [475.4220992391843, 0.0, 378.1094270899403],
    [0.0, 475.6638941565314, 236.6226272309063],
    [0.0, 0.0, 1.0],
], dtype=np.float64)
D_README = np.array([-0.2987656130625547, 0.090512327786479, 0.0006983134447049677,
                     0.0004069824038616868, 0.0], dtype=np.float64)

# README.md:338,356-361 ("default" = README launch values); names follow pose_estimator.h:139-155
DEFAULT_PARAMS = dict(
    N_Particle=100,
    maxAngularNoise=0.015, minAngularNoise=-0.015,
    maxTransitionNoise=0.035, minTransitionNoise=-0.035,
    back_projection_pixel_tolerance_PF=4.0,
    back_projection_pixel_tolerance=5.0,
    bMarkerDowngrade=[0, 0, 0, 0, 0],
)

FPS = 50.0
DT = 1.0 / FPS


def rot_x(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)


def rot_y(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)


def rot_z(a):
    c, s = np.cos(a), np.sin(a)
    return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)


def true_pose(t, phase=0.0):
    """Object->camera pose T_co(t) of the synthetic trajectory."""
    t = t + phase
    T = np.eye(4)
    T[:3, :3] = rot_z(0.4 * np.sin(0.5 * t)) @ rot_y(0.3 * np.sin(0.8 * t + 1.0)) @ rot_x(0.3 * np.sin(0.6 * t + 2.0))
    T[:3, 3] = [0.3 * np.sin(0.7 * t), 0.2 * np.sin(1.1 * t + 0.5), 2.0 + 0.5 * np.sin(0.4 * t)]
    return T


def project(K, T, markers):
    """Pinhole projection of (L,3) markers through pose T (pose_estimator.cpp:1017-1034 semantics)."""
    pc = markers @ T[:3, :3].T + T[:3, 3]
    p = pc @ np.asarray(K).T
    return p[:, :2] / p[:, 2:3]


def blobs_for_frame(frame, tracker=0, n_outliers=10, markers=MARKERS_5, K=K_README, phase=None,
                    noise_px=0.3, occlude=()):
    """Undistorted detections for one frame: true projections + N(0, noise^2) + uniform outliers, shuffled."""
    if phase is None:
        phase = tracker_phase(tracker)
    rng = np.random.default_rng(20240001 + frame + 1000003 * tracker)
    T = true_pose(frame * DT, phase)
    uv = project(K, T, markers) + rng.normal(0.0, noise_px, size=(markers.shape[0], 2))
    keep = [j for j in range(markers.shape[0]) if j not in occlude]
    uv = uv[keep]
    out = np.stack([rng.uniform(0, IMAGE_W, size=n_outliers), rng.uniform(0, IMAGE_H, size=n_outliers)], axis=1)
    pts = np.concatenate([uv, out], axis=0)
    rng.shuffle(pts, axis=0)
    return np.ascontiguousarray(pts)


def tracker_phase(tracker):
    return 0.37 * tracker


def pf_seed(frame, tracker=0):
    """Injected value of rd() for (frame, tracker): 1 + f + 100003*r."""
    return (1 + frame + 100003 * tracker) & 0xFFFFFFFF


class Scene:
    """One tracker's synthetic sequence: poses for frames 0/1 stand in for initialise();
    PF updates start at frame 2."""

    def __init__(self, tracker=0, n_outliers=10, markers=MARKERS_5, K=K_README, params=None):
        self.tracker = tracker
        self.n_outliers = n_outliers
        self.markers = np.asarray(markers, dtype=np.float64)
        self.K = np.asarray(K, dtype=np.float64)
        self.params = dict(DEFAULT_PARAMS if params is None else params)
        self.phase = tracker_phase(tracker)

    def pose(self, frame):
        return true_pose(frame * DT, self.phase)

    def time(self, frame):
        return frame * DT

    def blobs(self, frame, occlude=()):
        return blobs_for_frame(frame, self.tracker, self.n_outliers, self.markers, self.K, self.phase,
                               occlude=occlude)

    def seed(self, frame):
        return pf_seed(frame, self.tracker)


def blob_tensor(scenes, frames, d_max=None):
    """Pack blobs of several scenes/frames into (n_trackers, n_frames, d_max, 2) float64 + counts."""
    per = [[s.blobs(f) for f in frames] for s in scenes]
    if d_max is None:
        d_max = max(b.shape[0] for row in per for b in row)
    out = np.zeros((len(scenes), len(frames), d_max, 2), dtype=np.float64)
    cnt = np.zeros((len(scenes), len(frames)), dtype=np.int32)
    for i, row in enumerate(per):
        for j, b in enumerate(row):
            out[i, j, : b.shape[0]] = b
            cnt[i, j] = b.shape[0]
    return out, cnt


def true_pose_batch(t, phase):
    """Vectorised true_pose: t scalar, phase (T,) -> (T,4,4)."""
    tt = t + np.asarray(phase, dtype=np.float64)
    n = tt.shape[0]

    def rz(a):
        c, s = np.cos(a), np.sin(a)
        R = np.zeros((n, 3, 3)); R[:, 0, 0] = c; R[:, 0, 1] = -s; R[:, 1, 0] = s; R[:, 1, 1] = c; R[:, 2, 2] = 1
        return R

    def ry(a):
        c, s = np.cos(a), np.sin(a)
        R = np.zeros((n, 3, 3)); R[:, 0, 0] = c; R[:, 0, 2] = s; R[:, 1, 1] = 1; R[:, 2, 0] = -s; R[:, 2, 2] = c
        return R

    def rx(a):
        c, s = np.cos(a), np.sin(a)
        R = np.zeros((n, 3, 3)); R[:, 0, 0] = 1; R[:, 1, 1] = c; R[:, 1, 2] = -s; R[:, 2, 1] = s; R[:, 2, 2] = c
        return R

    T = np.zeros((n, 4, 4))
    T[:, :3, :3] = rz(0.4 * np.sin(0.5 * tt)) @ ry(0.3 * np.sin(0.8 * tt + 1.0)) @ rx(0.3 * np.sin(0.6 * tt + 2.0))
    T[:, 0, 3] = 0.3 * np.sin(0.7 * tt)
    T[:, 1, 3] = 0.2 * np.sin(1.1 * tt + 0.5)
    T[:, 2, 3] = 2.0 + 0.5 * np.sin(0.4 * tt)
    T[:, 3, 3] = 1.0
    return T


def blob_tensor_fast(trackers, frames, n_outliers=10, markers=MARKERS_5, K=K_README, noise_px=0.3, seed=20240001):
    """Vectorised synthetic detections for many trackers (bench workloads): same distribution as
    blobs_for_frame (true projections + N(0, noise^2) + uniform outliers, shuffled), one seeded
    generator for the whole tensor.  Returns (T, F, D, 2) float64 and (T, F) int32 counts."""
    trackers = np.asarray(trackers)
    phase = tracker_phase(trackers.astype(np.float64))
    rng = np.random.default_rng(seed)
    L = markers.shape[0]
    D = L + n_outliers
    out = np.zeros((len(trackers), len(frames), D, 2))
    for j, f in enumerate(frames):
        T = true_pose_batch(f * DT, phase)
        pc = np.einsum("tij,lj->tli", T[:, :3, :3], markers) + T[:, None, :3, 3]
        p = pc @ np.asarray(K).T
        uv = p[:, :, :2] / p[:, :, 2:3] + rng.normal(0.0, noise_px, size=(len(trackers), L, 2))
        o = np.stack([rng.uniform(0, IMAGE_W, size=(len(trackers), n_outliers)),
                      rng.uniform(0, IMAGE_H, size=(len(trackers), n_outliers))], axis=-1)
        pts = np.concatenate([uv, o], axis=1)
        perm = np.argsort(rng.uniform(size=(len(trackers), D)), axis=1)
        out[:, j] = np.take_along_axis(pts, perm[:, :, None], axis=1)
    cnt = np.full((len(trackers), len(frames)), D, dtype=np.int32)
    return out, cnt
