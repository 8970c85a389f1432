"""Seeded synthetic scenes of the shapes BASELINE.json names.

Geometry: cameras on an arc of a ring looking at a central landmark cloud, so that every (camera, landmark)
pair is geometrically visible and the track structure (who sees what) can be chosen freely:

* each landmark gets a track length  t = 2 + Geometric  (mean ``mean_track``), observed from cameras inside a
  window of ``window`` neighbouring cameras around a random centre camera; a fraction ``far_frac`` of the
  observations is re-drawn from ALL cameras (community-photo style co-visibility, K5);
* observations = exact pinhole projection + N(0, 0.5 px), 1 % gross outliers (uniform +-20 px);
* the initial guess is the truth perturbed (rotation N(0, 0.5 deg)/axis, translation 1 % of the scene scale,
  landmarks 1 % of the depth) and then ROUNDED THE WAY THE REFERENCE STORES IT: keypoints, intrinsics,
  landmarks and the camera->world 4x4 poses are float32 (Frame.h:87, MapPoint.h:69), camera 0 is the gauge
  anchor (frame ID 0 -> point-only residuals with its float32 extrinsics baked in, Optimizer.cpp:315-321).

Everything is numpy, vectorised (K5 = 30 M observations generates in well under a minute).
"""
from __future__ import annotations

import dataclasses
import numpy as np

# intrinsics hard-coded by the reference's sensors
TUM_K = (525.0, 525.0, 319.5, 239.5, 640, 480)        # VirtualSensor.h:154-162
REPLICA_K = (600.0, 600.0, 599.5, 339.5, 1200, 680)   # VirtualSensor.h:100-114


@dataclasses.dataclass
class Scene:
    """Reference-side view of a problem (what SceneMap/Frame/MapPoint hold)."""
    name: str
    seed: int
    pose_f32: np.ndarray      # [Nc,4,4] camera->world, float32 (Frame::getPose)
    K_f32: np.ndarray         # [Nc,3,3] float32 (Frame::getIntrinsics)
    frame_id: np.ndarray      # [Nc] int32; frame ID 0 is the anchor
    pt_f32: np.ndarray        # [Np,3] float32 (MapPoint::getPosition)
    obs_cam: np.ndarray       # [No] int32
    obs_pt: np.ndarray        # [No] int32
    obs_uv: np.ndarray        # [No,2] float32 (KeyPoint::pt)
    obs_octave: np.ndarray    # [No] int32 (KeyPoint::octave)
    width: int
    height: int
    gt_pose: np.ndarray       # [Nc,4,4] float64 ground truth camera->world
    gt_pt: np.ndarray         # [Np,3] float64


@dataclasses.dataclass
class Problem:
    """Flat SoA problem = the C-ABI input (include/ba_b200.h)."""
    cam: np.ndarray           # [Nc,6] f64 angle-axis + translation (world->camera)
    cam_fixed: np.ndarray     # [Nc] u8
    fixed_extr: np.ndarray    # [Nc,16] f32 row-major world->camera (read only where cam_fixed)
    K: np.ndarray             # [Nc,9] f32 row-major
    pt: np.ndarray            # [Np,3] f64
    obs_cam: np.ndarray       # [No] i32
    obs_pt: np.ndarray        # [No] i32
    obs_uv: np.ndarray        # [No,2] f32

    @property
    def shape(self):
        return (self.cam.shape[0], self.pt.shape[0], self.obs_cam.shape[0])


def _rodrigues(w: np.ndarray) -> np.ndarray:
    """[N,3] angle-axis -> [N,3,3] rotation (float64)."""
    th = np.linalg.norm(w, axis=1)
    small = th < 1e-12
    ths = np.where(small, 1.0, th)
    k = w / ths[:, None]
    Kx = np.zeros((w.shape[0], 3, 3))
    Kx[:, 0, 1], Kx[:, 0, 2] = -k[:, 2], k[:, 1]
    Kx[:, 1, 0], Kx[:, 1, 2] = k[:, 2], -k[:, 0]
    Kx[:, 2, 0], Kx[:, 2, 1] = -k[:, 1], k[:, 0]
    s, c = np.sin(th)[:, None, None], np.cos(th)[:, None, None]
    R = np.eye(3)[None] + s * Kx + (1 - c) * (Kx @ Kx)
    R[small] = np.eye(3)
    return R


def rotation_to_angle_axis(R: np.ndarray) -> np.ndarray:
    """[N,3,3] -> [N,3]; same quaternion route as ceres::RotationMatrixToAngleAxis (trace >= 0 branch is the
    only one synthetic scenes hit, the general case is handled for completeness)."""
    N = R.shape[0]
    q = np.zeros((N, 4))
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    pos = tr >= 0
    t = np.sqrt(np.where(pos, tr, 0) + 1.0)
    q[pos, 0] = 0.5 * t[pos]
    tt = 0.5 / t
    q[pos, 1] = ((R[:, 2, 1] - R[:, 1, 2]) * tt)[pos]
    q[pos, 2] = ((R[:, 0, 2] - R[:, 2, 0]) * tt)[pos]
    q[pos, 3] = ((R[:, 1, 0] - R[:, 0, 1]) * tt)[pos]
    for n in np.nonzero(~pos)[0]:
        Rn = R[n]
        i = 0
        if Rn[1, 1] > Rn[0, 0]:
            i = 1
        if Rn[2, 2] > Rn[i, i]:
            i = 2
        j, k = (i + 1) % 3, (i + 2) % 3
        t_ = np.sqrt(Rn[i, i] - Rn[j, j] - Rn[k, k] + 1.0)
        q[n, i + 1] = 0.5 * t_
        t_ = 0.5 / t_
        q[n, 0] = (Rn[k, j] - Rn[j, k]) * t_
        q[n, j + 1] = (Rn[j, i] + Rn[i, j]) * t_
        q[n, k + 1] = (Rn[k, i] + Rn[i, k]) * t_
    s2 = (q[:, 1:] ** 2).sum(1)
    st = np.sqrt(s2)
    ct = q[:, 0]
    two_theta = 2.0 * np.where(ct < 0, np.arctan2(-st, -ct), np.arctan2(st, ct))
    kk = np.where(s2 > 0, two_theta / np.where(s2 > 0, st, 1.0), 2.0)
    return q[:, 1:] * kk[:, None]


def make_scene(name: str, seed: int, n_cam: int, n_pt: int, mean_track: float, window: int,
               intr=TUM_K, far_frac: float = 0.0, arc_deg: float = 90.0, ring_radius: float = 4.0,
               cloud_radius: float = 1.0, noise_px: float = 0.5, outlier_frac: float = 0.01) -> Scene:
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy, W, H = intr
    # ---- ground-truth cameras: on an arc, looking at the origin, with a slow wobble ----
    i = np.arange(n_cam)
    phi = np.deg2rad(arc_deg) * (i / max(n_cam - 1, 1) - 0.5)
    rad = ring_radius * (1.0 + 0.05 * np.sin(3.0 * phi + 0.3))
    C = np.stack([rad * np.sin(phi), 0.15 * ring_radius * np.sin(2.0 * phi + 1.0), -rad * np.cos(phi)], 1)
    target = 0.1 * cloud_radius * np.stack([np.sin(5 * phi), np.cos(3 * phi), np.sin(2 * phi)], 1)
    z = target - C
    z /= np.linalg.norm(z, axis=1, keepdims=True)
    up = np.array([0.0, 1.0, 0.0])
    x = np.cross(up[None], z)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    y = np.cross(z, x)
    Rwc = np.stack([x, y, z], 2)                     # camera->world rotation (columns = camera axes)
    gt_pose = np.tile(np.eye(4), (n_cam, 1, 1))
    gt_pose[:, :3, :3] = Rwc
    gt_pose[:, :3, 3] = C
    # ---- ground-truth landmarks: uniform in a ball ----
    d = rng.normal(size=(n_pt, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    gt_pt = d * (cloud_radius * rng.random(n_pt) ** (1.0 / 3.0))[:, None]
    # ---- tracks ----
    extra_mean = max(mean_track - 2.0, 1e-3)
    t = 2 + rng.geometric(1.0 / (extra_mean + 1.0), size=n_pt) - 1
    win = 1 << int(np.floor(np.log2(max(min(window, n_cam), 1))))   # power of two so odd strides are coprime
    t = np.minimum(t, win)
    obs_pt = np.repeat(np.arange(n_pt, dtype=np.int64), t)
    n_obs0 = obs_pt.shape[0]
    centre = rng.integers(0, n_cam, size=n_pt)
    lo = np.clip(centre - win // 2, 0, max(n_cam - win, 0))
    # t distinct cameras inside the window: (start + k * odd_stride) mod win, k = 0..t-1
    start = rng.integers(0, win, size=n_pt)
    stride = 2 * rng.integers(0, max(win // 2, 1), size=n_pt) + 1
    first = np.cumsum(t) - t
    k = np.arange(n_obs0, dtype=np.int64) - np.repeat(first, t)
    off = (start[obs_pt] + k * stride[obs_pt]) % win
    obs_cam = lo[obs_pt] + off
    if far_frac > 0:
        far = rng.random(n_obs0) < far_frac
        obs_cam = np.where(far, rng.integers(0, n_cam, size=n_obs0), obs_cam)
    key = obs_pt * np.int64(n_cam) + obs_cam
    key.sort()                                                    # sorted by (pt, cam) ...
    if key.size:
        key = key[np.concatenate([[True], key[1:] != key[:-1]])]  # ... and deduped (== np.unique, 6x faster)
    obs_pt = (key // n_cam).astype(np.int64)
    obs_cam = (key % n_cam).astype(np.int64)
    # ---- project ----
    Rcw = np.transpose(Rwc, (0, 2, 1))
    tcw = -np.einsum('nij,nj->ni', Rcw, C)
    Xc = np.einsum('nij,nj->ni', Rcw[obs_cam], gt_pt[obs_pt]) + tcw[obs_cam]
    u = fx * Xc[:, 0] / Xc[:, 2] + cx
    v = fy * Xc[:, 1] / Xc[:, 2] + cy
    uv = np.stack([u, v], 1) + rng.normal(scale=noise_px, size=(key.shape[0], 2))
    # gross outliers only on landmarks with >= 3 views: a 2-view landmark with one wrong match is not
    # triangulable (the reference's pipeline culls those before BA, SfMHelper.cpp:974-1077)
    views = np.bincount(obs_pt, minlength=n_pt)
    gross = (rng.random(key.shape[0]) < outlier_frac) & (views[obs_pt] >= 3)
    uv = np.where(gross[:, None], uv + rng.uniform(-20.0, 20.0, size=uv.shape), uv)
    keep = (Xc[:, 2] > 0) & (u >= 0) & (u < W) & (v >= 0) & (v < H)
    # landmarks left with < 2 observations lose them all (they stay in the map, unobserved)
    cnt = np.bincount(obs_pt[keep], minlength=n_pt)
    keep &= cnt[obs_pt] >= 2
    obs_pt, obs_cam, uv = obs_pt[keep], obs_cam[keep], uv[keep]
    octave = rng.integers(0, 4, size=obs_pt.shape[0]).astype(np.int32)
    # ---- initial guess = perturbed truth, stored the way the reference stores it ----
    scene_scale = ring_radius
    dR = _rodrigues(rng.normal(scale=np.deg2rad(0.5), size=(n_cam, 3)))
    pose0 = gt_pose.copy()
    pose0[:, :3, :3] = np.einsum('nij,njk->nik', Rwc, dR)
    pose0[:, :3, 3] = C + rng.normal(scale=0.01 * scene_scale, size=(n_cam, 3))
    pose0[0] = gt_pose[0]                                       # the anchor frame keeps its pose
    depth = np.linalg.norm(gt_pt - C[n_cam // 2][None], axis=1)
    pt0 = gt_pt + rng.normal(size=(n_pt, 3)) * (0.01 * depth)[:, None]
    Kmat = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], dtype=np.float32)
    return Scene(name=name, seed=seed, pose_f32=pose0.astype(np.float32),
                 K_f32=np.tile(Kmat, (n_cam, 1, 1)), frame_id=np.arange(n_cam, dtype=np.int32),
                 pt_f32=pt0.astype(np.float32), obs_cam=obs_cam.astype(np.int32),
                 obs_pt=obs_pt.astype(np.int32), obs_uv=uv.astype(np.float32), obs_octave=octave,
                 width=W, height=H, gt_pose=gt_pose, gt_pt=gt_pt)


def flatten(scene: Scene) -> Problem:
    """numpy restatement of prepareConstraints' flattening (Optimizer.cpp:282-299): extr = inverse(pose) in
    float32, rotation -> angle-axis in double, translation promoted to double; frame ID 0 becomes a fixed
    camera whose float32 extrinsics are baked into point-only residuals.  (The product's own flattening is the
    C++ host adaptor / ba_pose_to_cam; this helper exists so CPU-only tests can build problems.)"""
    extr = np.linalg.inv(scene.pose_f32.astype(np.float32)).astype(np.float32)
    R = extr[:, :3, :3].astype(np.float64)
    cam = np.concatenate([rotation_to_angle_axis(R), extr[:, :3, 3].astype(np.float64)], 1)
    fixed = (scene.frame_id == 0).astype(np.uint8)
    return Problem(cam=np.ascontiguousarray(cam), cam_fixed=fixed,
                   fixed_extr=np.ascontiguousarray(extr.reshape(-1, 16)),
                   K=np.ascontiguousarray(scene.K_f32.reshape(-1, 9)),
                   pt=np.ascontiguousarray(scene.pt_f32.astype(np.float64)),
                   obs_cam=np.ascontiguousarray(scene.obs_cam), obs_pt=np.ascontiguousarray(scene.obs_pt),
                   obs_uv=np.ascontiguousarray(scene.obs_uv))


# ---- the five named configurations -------------------------------------------------
def config(name: str, scale: float = 1.0) -> Scene:
    """K1..K5; ``scale`` < 1 shrinks camera and landmark counts proportionally (parity-test sizes)."""
    s = lambda n: max(int(round(n * scale)), 4)
    if name == "K1":   # TUM fr1/xyz-shaped: 30 keyframes, 10k landmarks, ~60k observations
        return make_scene("K1", 1, s(30), s(10_000), 6.0, 16, TUM_K, arc_deg=60.0)
    if name == "K2":   # Replica-shaped 200 frames, 60k landmarks, ~300k observations
        return make_scene("K2", 2, s(200), s(60_000), 5.0, 16, REPLICA_K, arc_deg=150.0, ring_radius=3.5)
    if name == "K4":   # BAL-Ladybug scale: 1.7k cameras, 500k landmarks, 2M observations
        return make_scene("K4", 4, s(1_700), s(500_000), 4.0, 32, TUM_K, arc_deg=330.0, ring_radius=12.0,
                          cloud_radius=3.0)
    if name == "K5":   # BAL-Final scale: 13k cameras, 4.5M landmarks, 30M observations
        return make_scene("K5", 5, s(13_000), s(4_500_000), 6.67, 64, TUM_K, far_frac=0.10, arc_deg=350.0,
                          ring_radius=14.0, cloud_radius=3.5)
    raise ValueError(f"unknown config {name!r} (K3 is built by make_pose_batch)")


@dataclasses.dataclass
class PoseBatch:
    """K3: B independent frame-to-frame pose problems (MotionOnlyBAOptimizerAngles inputs)."""
    cam: np.ndarray        # [B,6] f64 initial angle-axis + translation
    K: np.ndarray          # [B,9] f32
    prob_ptr: np.ndarray   # [B+1] i32
    Xh: np.ndarray         # [No,4] f32 homogeneous landmark (MapPoint::getPosition().homogeneous())
    uv: np.ndarray         # [No,2] f32
    gt_cam: np.ndarray     # [B,6] f64


def make_pose_batch(seed: int = 3, n_prob: int = 1000, n_pts: int = 500, intr=TUM_K) -> PoseBatch:
    rng = np.random.default_rng(seed)
    fx, fy, cx, cy, W, H = intr
    B = n_prob
    w_gt = rng.normal(scale=0.05, size=(B, 3))
    t_gt = rng.normal(scale=0.1, size=(B, 3))
    R = _rodrigues(w_gt)
    # landmarks in front of the camera frame, transformed back to world
    Xc = np.stack([rng.uniform(-1.0, 1.0, (B, n_pts)), rng.uniform(-0.8, 0.8, (B, n_pts)),
                   rng.uniform(1.5, 5.0, (B, n_pts))], 2)
    Xw = np.einsum('bji,bnj->bni', R, Xc - t_gt[:, None, :])        # R^T (Xc - t)
    Xw32 = Xw.astype(np.float32)
    Xc32 = np.einsum('bij,bnj->bni', R, Xw32.astype(np.float64)) + t_gt[:, None, :]
    u = fx * Xc32[..., 0] / Xc32[..., 2] + cx
    v = fy * Xc32[..., 1] / Xc32[..., 2] + cy
    uv = np.stack([u, v], 2) + rng.normal(scale=0.5, size=(B, n_pts, 2))
    gross = rng.random((B, n_pts)) < 0.02
    uv = np.where(gross[..., None], uv + rng.uniform(-20, 20, size=uv.shape), uv)
    keep = (u >= 0) & (u < W) & (v >= 0) & (v < H)
    counts = keep.sum(1)
    prob_ptr = np.zeros(B + 1, dtype=np.int32)
    prob_ptr[1:] = np.cumsum(counts)
    Xh = np.concatenate([Xw32, np.ones((B, n_pts, 1), np.float32)], 2)[keep]
    uvk = uv[keep].astype(np.float32)
    cam0 = np.concatenate([w_gt + rng.normal(scale=np.deg2rad(1.0), size=(B, 3)),
                           t_gt + rng.normal(scale=0.03, size=(B, 3))], 1)
    Kmat = np.array([fx, 0, cx, 0, fy, cy, 0, 0, 1], dtype=np.float32)
    return PoseBatch(cam=np.ascontiguousarray(cam0), K=np.tile(Kmat, (B, 1)), prob_ptr=prob_ptr,
                     Xh=np.ascontiguousarray(Xh), uv=np.ascontiguousarray(uvk),
                     gt_cam=np.concatenate([w_gt, t_gt], 1))
