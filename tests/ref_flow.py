"""Oracle-side restatement of the reference's optimizer ENTRY POINTS (the outer loops around ceres::Solve):
GlobalBAOptimizerAngles::optimizeCamerasAndMapPoints (Optimizer.cpp:216-277), LocalBAOptimizerAngles (:500-610) and
MotionOnlyBAOptimizerAngles::optimizeCameraPose (:417-457), on flat "object graph" arrays.  TEST INFRASTRUCTURE: used to
check the C++ drop-in classes (bundleadjustment_b200/host/Optimizer.cpp) end to end."""
from __future__ import annotations

import ctypes as C
import dataclasses
import os
import subprocess

import numpy as np

from bundleadjustment_b200.synthetic import Problem, PoseBatch

HERE = os.path.dirname(os.path.abspath(__file__))
SHIM = os.path.join(HERE, "host", "_build", "libba_host_shim.so")


@dataclasses.dataclass
class Graph:
    """What SceneMap / Frame / MapPoint hold, flattened."""
    pose: np.ndarray        # [Nc,4,4] f32 camera->world
    K: np.ndarray           # [Nc,3,3] f32
    frame_id: np.ndarray    # [Nc] i32
    kp_ptr: np.ndarray      # [Nc+1] i32
    kp_xy: np.ndarray       # [Nk,2] f32
    kp_octave: np.ndarray   # [Nk] i32
    kp_point: np.ndarray    # [Nk] i32, -1 = no map point
    kp_outlier: np.ndarray  # [Nk] u8
    pt: np.ndarray          # [Np,3] f32
    pt_min: np.ndarray      # [Np] f32
    pt_max: np.ndarray      # [Np] f32


def graph_from_scene(sc, extra_kp=3, seed=0) -> Graph:
    """Frame f's keypoints = its observations (in scene order) + a few keypoints without a map point."""
    rng = np.random.default_rng(seed)
    Nc = sc.pose_f32.shape[0]
    order = np.argsort(sc.obs_cam, kind="stable")
    ptr, xy, octv, kpp = [0], [], [], []
    for f in range(Nc):
        sel = order[sc.obs_cam[order] == f]
        xy.append(sc.obs_uv[sel]); octv.append(sc.obs_octave[sel]); kpp.append(sc.obs_pt[sel].astype(np.int32))
        xy.append(rng.uniform(0, 400, size=(extra_kp, 2)).astype(np.float32))
        octv.append(np.zeros(extra_kp, np.int32)); kpp.append(np.full(extra_kp, -1, np.int32))
        ptr.append(ptr[-1] + sel.shape[0] + extra_kp)
    d = np.linalg.norm(sc.pt_f32 - sc.pose_f32[Nc // 2, :3, 3], axis=1)
    return Graph(pose=sc.pose_f32.copy(), K=sc.K_f32.copy(), frame_id=sc.frame_id.copy(), kp_ptr=np.array(ptr, np.int32),
                 kp_xy=np.concatenate(xy).astype(np.float32), kp_octave=np.concatenate(octv).astype(np.int32),
                 kp_point=np.concatenate(kpp), kp_outlier=np.zeros(ptr[-1], np.uint8), pt=sc.pt_f32.copy(),
                 pt_min=(0.2 * d).astype(np.float32), pt_max=(3.0 * d).astype(np.float32))


# --------------------------------------------------------------------------------------------------- the oracle flow
def _flatten(orc, g: Graph, free, fixed, points):
    """prepareConstraints (Optimizer.cpp:279-333 / 612-698) -> synthetic.Problem over cameras free+fixed."""
    cams = list(free) + list(fixed)
    pidx = {p: i for i, p in enumerate(points)}
    cam = np.zeros((len(cams), 6)); extr = np.zeros((len(cams), 16), np.float32)
    fixed_flag = np.zeros(len(cams), np.uint8)
    oc, op, uv = [], [], []
    for i, f in enumerate(cams):
        cam[i], e = orc.pose_to_cam(g.pose[f])
        extr[i] = e.reshape(16)
        fixed_flag[i] = 1 if (i >= len(free) or g.frame_id[f] == 0) else 0
        for k in range(g.kp_ptr[f], g.kp_ptr[f + 1]):
            p = int(g.kp_point[k])
            if p < 0 or g.kp_outlier[k] or p not in pidx:
                continue
            oc.append(i); op.append(pidx[p]); uv.append(g.kp_xy[k])
    return Problem(cam=cam, cam_fixed=fixed_flag, fixed_extr=extr, K=np.ascontiguousarray(g.K[cams].reshape(-1, 9)),
                   pt=np.ascontiguousarray(g.pt[points].astype(np.float64)).reshape(-1, 3), obs_cam=np.array(oc, np.int32),
                   obs_pt=np.array(op, np.int32), obs_uv=np.array(uv, np.float32).reshape(-1, 2))


def _prune(orc, g: Graph, frames, consider):
    for f in frames:
        a, b = g.kp_ptr[f], g.kp_ptr[f + 1]
        has = (g.kp_point[a:b] >= 0).astype(np.uint8)
        pi = np.where(g.kp_point[a:b] >= 0, g.kp_point[a:b], 0)
        g.kp_outlier[a:b] = orc.prune_frame(g.pose[f], g.K[f], g.kp_xy[a:b], g.kp_octave[a:b], g.pt[pi], g.pt_min[pi],
                                            g.pt_max[pi], has, g.kp_outlier[a:b], consider)


def _solve_and_write_back(orc, g, free, fixed, points, max_it):
    pb = _flatten(orc, g, free, fixed, points)
    cam, pt, sm, its = orc.solve(pb, orc.default_options(max_num_iterations=max_it, num_threads=8))
    g.pt[points] = pt.astype(np.float32)
    for i, f in enumerate(free):
        g.pose[f] = orc.cam_to_pose(cam[i])
    return sm


def global_ba(orc, g: Graph, n_outer=3, max_it=200) -> Graph:
    g = dataclasses.replace(g, pose=g.pose.copy(), pt=g.pt.copy(), kp_outlier=g.kp_outlier.copy())
    frames = list(np.argsort(g.frame_id, kind="stable"))
    points = list(range(g.pt.shape[0]))
    for _ in range(n_outer):
        g.sm = _solve_and_write_back(orc, g, frames, [], points, max_it)
        _prune(orc, g, frames, False)
    return g


def local_window(g: Graph, current, cov):
    """Optimizer.cpp:502-546."""
    local = list(cov[:10]) + [current]
    pts, seen = [], set()
    for f in local:
        for k in range(g.kp_ptr[f], g.kp_ptr[f + 1]):
            p = int(g.kp_point[k])
            if p >= 0 and not g.kp_outlier[k] and p not in seen:
                seen.add(p); pts.append(p)
    # observers of the local points, ordered like std::map<shared_ptr<Frame>>: by ADDRESS in the reference; the set of
    # fixed frames is what matters (their order only permutes residual blocks)
    fixed, fs = [], set(local)
    for p in pts:
        for f in range(g.pose.shape[0]):
            ks = np.nonzero(g.kp_point[g.kp_ptr[f]:g.kp_ptr[f + 1]] == p)[0]
            if ks.size and not g.kp_outlier[g.kp_ptr[f] + ks[0]] and f not in fs:
                fs.add(f); fixed.append(f)
    return local, fixed, pts


def local_ba(orc, g: Graph, current, cov, n_outer=1, max_it=10) -> Graph:
    g = dataclasses.replace(g, pose=g.pose.copy(), pt=g.pt.copy(), kp_outlier=g.kp_outlier.copy())
    local, fixed, pts = local_window(g, current, cov)
    for _ in range(n_outer):
        g.sm = _solve_and_write_back(orc, g, local, fixed, pts, max_it)
        _prune(orc, g, local + fixed, False)
    g.window = (local, fixed, pts)
    return g


def pose_graph_from_batch(batch: PoseBatch, seed=0) -> Graph:
    """K3's 1 000 independent frame-to-frame problems as an object graph: frame b observes its own private landmarks
    (no sharing between frames), the initial pose is the float32 camera->world matrix of batch.cam, a few keypoints start as
    outliers (they can become inliers again: considerOutlier = true) and a few have no map point."""
    rng = np.random.default_rng(seed)
    from oracle import ba_oracle as orc
    B = batch.cam.shape[0]
    n = int(batch.prob_ptr[-1])
    pose = np.stack([orc.cam_to_pose(batch.cam[b]) for b in range(B)]).astype(np.float32)
    pt = np.ascontiguousarray(batch.Xh[:, :3], np.float32)
    kp_frame = np.repeat(np.arange(B), np.diff(batch.prob_ptr))
    centre = pose[kp_frame, :3, 3]
    d = np.linalg.norm(pt - centre, axis=1)
    kp_point = np.arange(n, dtype=np.int32)
    kp_point[rng.random(n) < 0.01] = -1
    return Graph(pose=pose, K=np.ascontiguousarray(batch.K.reshape(B, 3, 3)), frame_id=np.arange(1, B + 1, dtype=np.int32),
                 kp_ptr=batch.prob_ptr.astype(np.int32), kp_xy=batch.uv.copy(), kp_octave=rng.integers(0, 4, n).astype(np.int32),
                 kp_point=kp_point, kp_outlier=(rng.random(n) < 0.03).astype(np.uint8), pt=pt,
                 pt_min=(0.2 * d).astype(np.float32), pt_max=(3.0 * d).astype(np.float32))


def keypoint_arrays(g: Graph):
    """per-keypoint map point data as ba_prune_frames / ba_solve_pose_only_outer take it"""
    has = (g.kp_point >= 0).astype(np.uint8)
    pi = np.where(g.kp_point >= 0, g.kp_point, 0)
    return g.pt[pi], g.pt_min[pi], g.pt_max[pi], has


def run_pose_outer(bab, g: Graph, n_outer=4, max_it=20) -> Graph:
    """the device-resident outer loop (ba_solve_pose_only_outer) on the same graph"""
    xyz, mn, mx, has = keypoint_arrays(g)
    pose, out, sms, ms, lm = bab.capi.solve_pose_only_outer(g.pose, g.K.reshape(-1, 9), g.kp_ptr, g.kp_xy, g.kp_octave, xyz, mn, mx, has,
                                                        g.kp_outlier, n_outer=n_outer, options=bab.default_options(max_num_iterations=max_it))
    res = dataclasses.replace(g, pose=pose.copy(), kp_outlier=out.copy())
    res.sms, res.device_ms, res.lm_iterations = sms, ms, lm
    return res


def motion_only(orc, g: Graph, n_outer=4, max_it=20, threads=8) -> Graph:
    g = dataclasses.replace(g, pose=g.pose.copy(), pt=g.pt.copy(), kp_outlier=g.kp_outlier.copy())
    Nc = g.pose.shape[0]
    for _ in range(n_outer):
        cam = np.stack([orc.pose_to_cam(g.pose[f])[0] for f in range(Nc)])
        keep = (g.kp_point >= 0) & (g.kp_outlier == 0)                      # Optimizer.cpp:472, keypoint order kept
        kp_frame = np.repeat(np.arange(Nc), np.diff(g.kp_ptr))
        ptr = np.zeros(Nc + 1, np.int32)
        ptr[1:] = np.cumsum(np.bincount(kp_frame[keep], minlength=Nc))
        Xh = np.concatenate([g.pt[g.kp_point[keep]], np.ones((int(keep.sum()), 1), np.float32)], 1).astype(np.float32)
        batch = PoseBatch(cam=cam, K=np.ascontiguousarray(g.K.reshape(-1, 9)), prob_ptr=ptr,
                          Xh=np.ascontiguousarray(Xh), uv=np.ascontiguousarray(g.kp_xy[keep], np.float32), gt_cam=cam)
        out, sms, _ = orc.solve_pose_only_batch(batch, orc.default_options(max_num_iterations=max_it, num_threads=threads))
        for f in range(Nc):
            g.pose[f] = orc.cam_to_pose(out[f])
        _prune(orc, g, range(Nc), True)
    return g


# --------------------------------------------------------------------------------------------------- the C++ drop-in
def build_shim():
    subprocess.check_call(["make", "-C", os.path.join(HERE, "host"), "-s"])
    return SHIM


def run_dropin(g: Graph, mode, n_outer, max_it, erase=False, current=-1, cov=()):
    lib = C.CDLL(build_shim())
    Nc, Np, Nk = g.pose.shape[0], g.pt.shape[0], g.kp_xy.shape[0]
    pose_out = np.zeros((Nc, 4, 4), np.float32); pt_out = np.zeros((Np, 3), np.float32)
    out_out = np.zeros(Nk, np.uint8); kpp_out = np.zeros(Nk, np.int32); alive = np.zeros(Np, np.uint8)
    info = np.zeros(4)
    cov = np.ascontiguousarray(cov, np.int32)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    arrs = [np.ascontiguousarray(x) for x in (g.pose, g.K, g.frame_id, g.kp_ptr, g.kp_xy, g.kp_octave, g.kp_point, g.kp_outlier)]
    pts = [np.ascontiguousarray(x) for x in (g.pt, g.pt_min, g.pt_max)]
    rc = lib.ba_host_shim_run(int(mode), Nc, *[P(a) for a in arrs], Np, *[P(a) for a in pts], int(n_outer), int(max_it),
                              int(erase), int(current), P(cov), int(cov.shape[0]), P(pose_out), P(pt_out), P(out_out),
                              P(kpp_out), P(alive), P(info))
    res = dataclasses.replace(g, pose=pose_out, pt=pt_out, kp_outlier=out_out, kp_point=kpp_out)
    res.alive = alive
    res.info = dict(status=rc, termination=int(info[0]), initial_cost=info[1], final_cost=info[2], num_iterations=int(info[3]))
    return res
