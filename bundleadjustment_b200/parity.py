"""Parity gate between two solutions of the SAME problem (host-side numpy; used by bench.py before any timing counts and by
tools/mgpu_check.py): sharded against single-GPU, or the CUDA path against the CPU baseline's solution.

Tolerances are north_star's: per-iteration cost trajectory 1e-6 relative, final cost 1e-8 relative, poses 1e-6, landmarks

    |x_a - x_b|  <=  1e-6 * max(1, |x|)  +  cond(V_p') * (C_COND * eps * |x_final - x_initial| + C_CAM * max|cam_a - cam_b|)

The second term is a STATED bound, not a waiver: a landmark seen along nearly parallel rays has an almost singular 3x3 block
V_p' = sum Jp^T Jp + D_p^2 (depth is barely observable), and two correct solvers that add the same terms in different orders
(landmark shards, tile plans, CPU vs GPU reductions) differ there by cond(V_p') times the rounding of their inputs, times the
distance the landmark moved, plus the response of the back-substitution y_p = V_p'^-1 (b_p - W^T y_c) to the (gated,
<= 1e-6) difference of the camera steps.  cond(V_p') is evaluated here, for the landmarks that need it only, from the pinhole Jacobians
of the reference functors (ba_project/src/ba/Optimizer.h:54-76, 96-107) at solution b; a landmark outside the bound fails
the gate."""
from __future__ import annotations

import numpy as np

EPS = float(np.finfo(np.float64).eps)
C_COND = 64.0
C_CAM = 8.0


def _rodrigues(w: np.ndarray) -> np.ndarray:
    th2 = (w * w).sum(1)
    big = th2 > EPS
    th = np.sqrt(np.where(big, th2, 1.0))
    k = w / th[:, None]
    c, s = np.cos(th), np.sin(th)
    R = np.einsum('ni,nj->nij', k, k) * (1 - c)[:, None, None]
    R[:, 0, 0] += c; R[:, 1, 1] += c; R[:, 2, 2] += c
    R[:, 0, 1] -= k[:, 2] * s; R[:, 0, 2] += k[:, 1] * s
    R[:, 1, 0] += k[:, 2] * s; R[:, 1, 2] -= k[:, 0] * s
    R[:, 2, 0] -= k[:, 1] * s; R[:, 2, 1] += k[:, 0] * s
    sm = ~big
    if sm.any():
        R[sm] = np.eye(3)
        R[sm, 0, 1] = -w[sm, 2]; R[sm, 0, 2] = w[sm, 1]; R[sm, 1, 0] = w[sm, 2]
        R[sm, 1, 2] = -w[sm, 0]; R[sm, 2, 0] = -w[sm, 1]; R[sm, 2, 1] = w[sm, 0]
    return R


def landmark_condition(pb, cam: np.ndarray, pt: np.ndarray, landmarks: np.ndarray, radius: float) -> np.ndarray:
    """cond_2(sum_o Jp_o^T Jp_o + diag(clamp(.)) / radius) for the given landmarks (unrobustified, unscaled Jacobians)."""
    landmarks = np.asarray(landmarks)
    pos = np.full(pb.pt.shape[0], -1, np.int64)
    pos[landmarks] = np.arange(landmarks.size)
    sel = np.nonzero(pos[pb.obs_pt] >= 0)[0]
    c, p = pb.obs_cam[sel], pb.obs_pt[sel]
    X = pt[p]
    K = pb.K.reshape(-1, 3, 3).astype(np.float64)[c]
    fixed = pb.cam_fixed[c] != 0
    R = _rodrigues(cam[:, :3])[c]
    Pm = np.concatenate([R, cam[c, 3:, None]], 2)                       # 3 x 4 [R | t]
    E = pb.fixed_extr.reshape(-1, 4, 4).astype(np.float64)[c]
    Pm = np.where(fixed[:, None, None], E[:, :3, :] / np.where(fixed, E[:, 3, 3], 1.0)[:, None, None], Pm)   # (affine extrinsics)
    q = np.einsum('nij,nj->ni', K, np.einsum('nij,nj->ni', Pm[:, :, :3], X) + Pm[:, :, 3])
    pix = q[:, :2] / q[:, 2:3]
    G = (K[:, :2, :] - pix[:, :, None] * K[:, 2:3, :]) / q[:, 2, None, None]      # d pix / d p_cam
    Jp = np.einsum('nij,njk->nik', G, Pm[:, :, :3])
    V = np.zeros((landmarks.size, 3, 3))
    np.add.at(V, pos[p], np.einsum('nki,nkj->nij', Jp, Jp))
    d = np.clip(np.einsum('nii->ni', V), 1e-6, 1e32) / radius
    V[:, 0, 0] += d[:, 0]; V[:, 1, 1] += d[:, 1]; V[:, 2, 2] += d[:, 2]
    w = np.linalg.eigvalsh(V)
    return w[:, 2] / np.maximum(w[:, 0], 1e-300)


def compare(pb, a: dict, b: dict, cost_rtol: float = 1e-6, final_rtol: float = 1e-8) -> dict:
    """a, b: {"costs": per-iteration costs, "final_cost", "cam" [Nc,6], "pt" [Np,3], "radius": last trust-region radius}.
    b is the side the bound is evaluated at (the single-GPU / CPU solution).  Returns the `parity` object of bench.py."""
    ca, cb = np.asarray(a["costs"], float), np.asarray(b["costs"], float)
    same_len = ca.shape == cb.shape
    cost_rel = float(np.max(np.abs(ca - cb) / np.abs(cb))) if same_len and cb.size else None
    final_rel = abs(a["final_cost"] - b["final_cost"]) / abs(b["final_cost"]) if b["final_cost"] else 0.0
    dcam = float(np.abs(a["cam"] - b["cam"]).max()) if a["cam"].size else 0.0
    diff = np.abs(a["pt"] - b["pt"]).max(axis=1) if a["pt"].size else np.zeros(0)
    rel = 1e-6 * np.maximum(1.0, np.abs(b["pt"]).max(axis=1)) if diff.size else diff
    need = np.nonzero(diff > rel)[0]
    n_bad = 0
    worst_ratio = 0.0
    if need.size:
        cond = landmark_condition(pb, b["cam"], b["pt"], need, float(b.get("radius", 1e4)))
        travel = np.linalg.norm(b["pt"][need] - pb.pt[need], axis=1)
        bound = rel[need] + cond * (C_COND * EPS * travel + C_CAM * dcam)
        n_bad = int((diff[need] > bound).sum())
        worst_ratio = float((diff[need] / bound).max())
    ok = bool(same_len and cost_rel is not None and cost_rel <= cost_rtol and final_rel <= final_rtol and dcam <= 1e-6 and n_bad == 0)
    return {"ok": ok, "iterations_equal": bool(same_len), "max_cost_rel_diff": cost_rel, "final_cost_rel_diff": float(final_rel),
            "max_cam_diff": dcam, "max_pt_diff": float(diff.max()) if diff.size else 0.0,
            "pt_over_1e-6_rel": int(need.size), "pt_outside_cond_bound": n_bad, "pt_worst_diff_over_bound": worst_ratio,
            "tolerances": {"cost_trajectory_rel": cost_rtol, "final_cost_rel": final_rtol, "cam_abs": 1e-6,
                           "pt": f"1e-6*max(1,|x|) + cond(V_p')*({C_COND:g}*eps*|x - x0| + {C_CAM:g}*max_cam_diff)"}}
