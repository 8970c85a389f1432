"""On-disk formats either side of the BA hot path, Python side: the same three formats as
host/ProblemIO.{h,cpp} (the C++ implementation a maintainer links), on `synthetic.Problem`:

* BAL text ("bundle adjustment in the large"): `cams pts obs` header, `cam pt x y` per observation, 9 numbers per camera
  (angle-axis, translation, f, k1, k2), 3 per point.  BAL's pinhole part (p = -P/P.z, pixel = f p) is exactly the
  reference's model (Optimizer.h:64-74) with K = diag(-f, -f, 1); k1/k2 are carried, not applied.
* flat binary dump "BAB200P\\1" of the ba_set_problem arrays (checkpoint / fixture format);
* the trajectory file of ba_project/src/ba/BundleAdjustment.cpp:248-268 (`timestamp tx ty tz qx qy qz qw`, 4 decimals).

Files written by either side are read by the other (tests/test_problem_io.py)."""
from __future__ import annotations

import numpy as np

from .synthetic import Problem, _rodrigues

MAGIC = b"BAB200P\x01"


def read_bal(path, fix_first_camera=True):
    """-> (Problem, distortion [Nc,2])."""
    with open(path, "rb") as f:
        tok = f.read().split()
    Nc, Np, No = int(tok[0]), int(tok[1]), int(tok[2])
    need = 3 + 4 * No + 9 * Nc + 3 * Np
    if len(tok) < need:
        raise ValueError(f"{path}: truncated ({len(tok)} tokens, {need} expected)")
    obs = np.array(tok[3:3 + 4 * No], dtype=np.float64).reshape(No, 4)
    obs_cam, obs_pt = obs[:, 0].astype(np.int32), obs[:, 1].astype(np.int32)
    if No and (obs_cam.min() < 0 or obs_cam.max() >= Nc or obs_pt.min() < 0 or obs_pt.max() >= Np):
        raise ValueError(f"{path}: observation index out of range")
    cams = np.array(tok[3 + 4 * No:3 + 4 * No + 9 * Nc], dtype=np.float64).reshape(Nc, 9)
    pt = np.array(tok[3 + 4 * No + 9 * Nc:need], dtype=np.float64).reshape(Np, 3)
    K = np.zeros((Nc, 9), np.float32)
    K[:, 0] = K[:, 4] = (-cams[:, 6]).astype(np.float32)
    K[:, 8] = 1.0
    cam_fixed = np.zeros(Nc, np.uint8)
    extr = np.zeros((Nc, 16), np.float32)
    if fix_first_camera and Nc:
        E = np.eye(4)
        E[:3, :3] = _rodrigues(cams[:1, :3])[0]
        E[:3, 3] = cams[0, 3:6]
        extr[0] = E.astype(np.float32).reshape(16)
        cam_fixed[0] = 1
    pb = Problem(cam=np.ascontiguousarray(cams[:, :6]), cam_fixed=cam_fixed, fixed_extr=extr, K=K, pt=pt, obs_cam=obs_cam,
                 obs_pt=obs_pt, obs_uv=np.ascontiguousarray(obs[:, 2:4].astype(np.float32)))
    return pb, np.ascontiguousarray(cams[:, 7:9])


def write_bal(path, pb: Problem, distortion=None):
    Nc, Np, No = pb.shape
    K = pb.K.reshape(Nc, 9)
    if not (np.all(K[:, 0] == K[:, 4]) and np.all(K[:, [1, 3, 6, 7]] == 0) and np.all(K[:, 8] == 1) and np.all(K[:, 0] != 0)):
        raise ValueError("K is not [[f,0,cx],[0,f,cy],[0,0,1]]; BAL cannot express it")
    s = np.where(K[pb.obs_cam, 0] < 0, 1.0, -1.0)
    uv = s[:, None] * (pb.obs_uv.astype(np.float64) - K[pb.obs_cam][:, [2, 5]].astype(np.float64))
    dist = np.zeros((Nc, 2)) if distortion is None else np.asarray(distortion, np.float64)
    with open(path, "w") as f:
        f.write(f"{Nc} {Np} {No}\n")
        f.writelines(f"{c} {p} {u:.9g} {v:.9g}\n" for c, p, (u, v) in zip(pb.obs_cam.tolist(), pb.obs_pt.tolist(), uv.tolist()))
        for c in range(Nc):
            f.writelines(f"{x:.17g}\n" for x in pb.cam[c].tolist())
            f.write(f"{abs(float(K[c, 0])):.9g}\n{dist[c, 0]:.17g}\n{dist[c, 1]:.17g}\n")
        f.writelines(f"{x:.17g}\n" for x in pb.pt.reshape(-1).tolist())


def _pad8(b: bytes) -> bytes:
    return b + b"\0" * ((8 - len(b) % 8) % 8)


def write_binary(path, pb: Problem, distortion=None):
    Nc, Np, No = pb.shape
    arrays = [(pb.cam, np.float64), (pb.cam_fixed, np.uint8), (pb.fixed_extr, np.float32), (pb.K, np.float32), (pb.pt, np.float64),
              (pb.obs_cam, np.int32), (pb.obs_pt, np.int32), (pb.obs_uv, np.float32)]
    with open(path, "wb") as f:
        f.write(MAGIC)
        f.write(np.array([Nc, Np, No, 0 if distortion is None else 1], "<i4").tobytes())
        for a, dt in arrays:
            f.write(_pad8(np.ascontiguousarray(a, dtype=np.dtype(dt).newbyteorder("<")).tobytes()))
        if distortion is not None:
            f.write(_pad8(np.ascontiguousarray(distortion, dtype="<f8").tobytes()))


def read_binary(path):
    """-> (Problem, distortion or None)."""
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:8] != MAGIC:
        raise ValueError(f"{path}: not a BAB200P file (or another version)")
    Nc, Np, No, has_d = np.frombuffer(buf, "<i4", 4, 8).tolist()
    off = 24

    def take(n, dt, shape):
        nonlocal off
        nbytes = n * np.dtype(dt).itemsize
        if off + nbytes > len(buf):
            raise ValueError(f"{path}: truncated")
        a = np.frombuffer(buf, dt, n, off).reshape(shape).copy()
        off += nbytes + (8 - nbytes % 8) % 8
        return a

    pb = Problem(cam=take(6 * Nc, "<f8", (Nc, 6)), cam_fixed=take(Nc, "u1", (Nc,)), fixed_extr=take(16 * Nc, "<f4", (Nc, 16)),
                 K=take(9 * Nc, "<f4", (Nc, 9)), pt=take(3 * Np, "<f8", (Np, 3)), obs_cam=take(No, "<i4", (No,)),
                 obs_pt=take(No, "<i4", (No,)), obs_uv=take(2 * No, "<f4", (No, 2)))
    dist = take(2 * Nc, "<f8", (Nc, 2)) if has_d else None
    if No and (pb.obs_cam.min() < 0 or pb.obs_cam.max() >= Nc or pb.obs_pt.min() < 0 or pb.obs_pt.max() >= Np):
        raise ValueError(f"{path}: observation index out of range")
    return pb, dist


def rotation_to_quaternion(R):
    """float32 3x3 -> (x, y, z, w), branch structure of Eigen::Quaternionf(Matrix3f) (BundleAdjustment.cpp:261)."""
    m = np.asarray(R, np.float32)
    q = np.zeros(4, np.float32)
    one, half = np.float32(1), np.float32(0.5)
    t = m[0, 0] + m[1, 1] + m[2, 2]
    if t > 0:
        t = np.sqrt(t + one)
        q[3] = half * t
        t = half / t
        q[0], q[1], q[2] = (m[2, 1] - m[1, 2]) * t, (m[0, 2] - m[2, 0]) * t, (m[1, 0] - m[0, 1]) * t
    else:
        i = 0
        if m[1, 1] > m[0, 0]:
            i = 1
        if m[2, 2] > m[i, i]:
            i = 2
        j = (i + 1) % 3
        k = (j + 1) % 3
        t = np.sqrt(m[i, i] - m[j, j] - m[k, k] + one)
        q[i] = half * t
        t = half / t
        q[3], q[j], q[k] = (m[k, j] - m[j, k]) * t, (m[j, i] + m[i, j]) * t, (m[k, i] + m[i, k]) * t
    return q


def write_trajectory_tum(output_path, timestamps, poses):
    """poses: [n,4,4] float32 camera->world, keyframe order.  Writes `<output_path>_estimatedPoses.txt`."""
    poses = np.asarray(poses, np.float32).reshape(-1, 4, 4)
    with open(str(output_path) + "_estimatedPoses.txt", "w") as f:
        for ts, P in zip(timestamps, poses):
            q = rotation_to_quaternion(P[:3, :3])
            vals = [float(ts)] + [float(x) for x in P[:3, 3]] + [float(x) for x in q]
            f.write(" ".join(f"{v:.4f}" for v in vals) + "\n")
