"""Independent pin of the OPTIMUM: scipy.optimize.least_squares on the reference's objective.

Neither the C++ oracle (oracle/ba_oracle.cpp) nor the CUDA path shares any code with this script or with scipy's
trust-region-reflective minimiser.  The objective is the one the reference hands to Ceres
(ba_project/src/ba/Optimizer.cpp:312-328): one residual block per observation, AngleReprojectionError
(Optimizer.h:54-76) for free cameras, PointOnlyReprojectionError (Optimizer.h:96-107, float32 extrinsics baked in) for
fixed ones, HuberLoss(sqrt(5.991)) on the block's squared norm s_i = ||res_i||^2:  F(x) = 0.5 * sum_i rho(s_i),
rho(s) = s (s <= a^2) else 2 a sqrt(s) - a^2.  scipy's built-in losses act per SCALAR residual, which is a different
objective for 2-vector blocks, so the blocks enter pre-robustified instead:  f_i = res_i * sqrt(rho(s_i) / s_i)  (a 2-vector
with ||f_i||^2 = rho(s_i)) under loss='linear'; scipy then minimises exactly F.  (The scalar form f_i = ||res_i|| with
loss='huber' is the same objective too, but its Gauss-Newton model is rank-1 per block and needed > 3000 evaluations.)
This is NOT Ceres' Corrector (which scales residual and Jacobian by sqrt(rho')), so the route to the optimum shares nothing
with the oracle's or the CUDA path's.

What this pins: cost semantics (pinhole + angle-axis + float32 promotion points), Huber, the handling of fixed cameras and
the minimiser the LM loop converges to.  What it cannot pin: the per-iteration trajectory (step acceptance, radius
schedule, Jacobi scaling) -- those remain anchored on Ceres' published algorithm only.

Monocular BA with ONE fixed camera keeps a scale gauge freedom (the optimum is a 1-D family), so parameters of two
different minimisers only agree in cost there.  Every case is therefore ALSO run with the second camera fixed as well
(suffix "_g2"): a unique local optimum, where cost (1e-8 relative) and parameters (1e-6) must both agree.

Run from the repo root:   python tests/golden/make_scipy_pin.py        (writes tests/golden/scipy_*.npz; ~2 minutes)
"""
import dataclasses
import os
import sys
import time

import numpy as np
from scipy.optimize import least_squares
from scipy.sparse import lil_matrix

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from bundleadjustment_b200 import synthetic as syn   # noqa: E402  (problem generator only)

A_HUBER = float(np.sqrt(5.991))
EPS = np.finfo(np.float64).eps


def rodrigues_batch(w):
    """ceres::AngleAxisToRotationMatrix for an [n,3] array (complex-safe: jac='cs')."""
    th2 = (w * w).sum(1)
    big = np.real(th2) > EPS
    th = np.sqrt(np.where(big, th2, 1.0))
    k = w / th[:, None]
    c, s = np.cos(th), np.sin(th)
    R = np.empty((w.shape[0], 3, 3), dtype=w.dtype)
    for i in range(3):
        for j in range(3):
            R[:, i, j] = k[:, i] * k[:, j] * (1 - c) + (c if i == j else 0)
    R[:, 0, 1] -= k[:, 2] * s; R[:, 0, 2] += k[:, 1] * s
    R[:, 1, 0] += k[:, 2] * s; R[:, 1, 2] -= k[:, 0] * s
    R[:, 2, 0] -= k[:, 1] * s; R[:, 2, 1] += k[:, 0] * s
    if not big.all():                              # first-order branch of ceres (theta^2 <= eps)
        sm = ~big
        one = np.ones(sm.sum(), dtype=w.dtype)
        R[sm] = np.stack([np.stack([one, -w[sm, 2], w[sm, 1]], 1), np.stack([w[sm, 2], one, -w[sm, 0]], 1),
                          np.stack([-w[sm, 1], w[sm, 0], one], 1)], 1)
    return R


class Objective:
    def __init__(self, pb, use_huber):
        self.pb = pb
        self.use_huber = use_huber
        Nc, Np, No = pb.shape
        self.free_cam = np.nonzero((pb.cam_fixed == 0) & (np.bincount(pb.obs_cam, minlength=Nc) > 0))[0]
        self.act_pt = np.nonzero(np.bincount(pb.obs_pt, minlength=Np) > 0)[0]
        self.cam_slot = np.full(Nc, -1); self.cam_slot[self.free_cam] = np.arange(self.free_cam.size)
        self.pt_slot = np.full(Np, -1); self.pt_slot[self.act_pt] = np.arange(self.act_pt.size)
        self.n = 6 * self.free_cam.size + 3 * self.act_pt.size
        self.K = pb.K.reshape(Nc, 3, 3).astype(np.float64)             # float32 promoted to T (Optimizer.h:70)
        self.E = pb.fixed_extr.reshape(Nc, 4, 4).astype(np.float64)    # float32 promoted to T (Optimizer.h:99)
        self.uv = pb.obs_uv.astype(np.float64)
        self.is_fixed = pb.cam_fixed[pb.obs_cam] != 0

    def pack(self, cam, pt):
        return np.concatenate([cam[self.free_cam].ravel(), pt[self.act_pt].ravel()])

    def unpack(self, x, cam0, pt0):
        cam, pt = cam0.astype(x.dtype).copy(), pt0.astype(x.dtype).copy()
        nc = 6 * self.free_cam.size
        cam[self.free_cam] = x[:nc].reshape(-1, 6)
        pt[self.act_pt] = x[nc:].reshape(-1, 3)
        return cam, pt

    def block_residuals(self, x):
        pb = self.pb
        cam, pt = self.unpack(x, pb.cam, pb.pt)
        X = pt[pb.obs_pt]
        c = pb.obs_cam
        R = rodrigues_batch(cam[:, :3])
        p_free = np.einsum('nij,nj->ni', R[c], X) + cam[c, 3:]
        Xh = np.concatenate([X, np.ones((X.shape[0], 1), dtype=X.dtype)], 1)
        ph = np.einsum('nij,nj->ni', self.E[c], Xh)
        p_fix = ph[:, :3] / ph[:, 3:4]                                  # (E X).hnormalized()
        p = np.where(self.is_fixed[:, None], p_fix, p_free)
        q = np.einsum('nij,nj->ni', self.K[c], p)
        return q[:, :2] / q[:, 2:3] - self.uv

    def fun(self, x):
        r = self.block_residuals(x)
        if self.use_huber:
            s = (r * r).sum(1)
            b = A_HUBER * A_HUBER
            out = np.real(s) > b
            ss = np.where(out, s, b)                                    # (inliers: any positive stand-in, masked below)
            r = r * np.where(out, np.sqrt((2 * A_HUBER * np.sqrt(ss) - b) / ss), 1.0)[:, None]
        return r.ravel()                                                # ||f_i||^2 = rho(s_i)

    def cost(self, x):
        s = (self.block_residuals(x) ** 2).sum(1)
        if self.use_huber:
            b = A_HUBER * A_HUBER
            s = np.where(s > b, 2 * A_HUBER * np.sqrt(s) - b, s)
        return 0.5 * float(s.sum())

    def sparsity(self):
        pb = self.pb
        S = lil_matrix((2 * pb.obs_cam.shape[0], self.n), dtype=np.int8)
        nc = 6 * self.free_cam.size
        for i, (c, p) in enumerate(zip(pb.obs_cam, pb.obs_pt)):
            for row in (2 * i, 2 * i + 1):
                if self.cam_slot[c] >= 0:
                    S[row, 6 * self.cam_slot[c]:6 * self.cam_slot[c] + 6] = 1
                S[row, nc + 3 * self.pt_slot[p]:nc + 3 * self.pt_slot[p] + 3] = 1
        return S


def fix_second_camera(pb):
    fixed = pb.cam_fixed.copy()
    free = np.nonzero(fixed == 0)[0]
    fixed[free[0]] = 1                     # its float32 extrinsics (flatten() filled fixed_extr for every camera) are baked in
    return dataclasses.replace(pb, cam_fixed=fixed)


def minimise(pb, use_huber, label):
    ob = Objective(pb, use_huber)
    x0 = ob.pack(pb.cam, pb.pt)
    t = time.time()
    big = ob.n > 400
    kw = dict(jac_sparsity=ob.sparsity(), tr_solver='lsmr', tr_options={"atol": 1e-14, "btol": 1e-14, "maxiter": 20000}) if big else {}
    res = least_squares(ob.fun, x0, jac='cs', loss='linear', method='trf', x_scale='jac', ftol=1e-15, xtol=1e-15, gtol=1e-11,
                        max_nfev=500, **kw)
    cam, pt = ob.unpack(res.x, pb.cam, pb.pt)
    cost = ob.cost(res.x)
    assert abs(cost - res.cost) <= 1e-12 * cost, (cost, res.cost)       # scipy's own value of the objective
    print(f"{label}: n={ob.n} cost {ob.cost(x0):.9e} -> {cost:.12e} nfev={res.nfev} status={res.status} "
          f"optimality={res.optimality:.2e} ({time.time() - t:.1f}s)")
    return cam, pt, cost, res


CASES = {
    # name: (scene builder, use_huber)
    "tiny_huber": (lambda: syn.make_scene("tiny_huber", 101, 4, 24, 3.0, 4, arc_deg=40.0), 1),
    "tiny_l2": (lambda: syn.make_scene("tiny_l2", 102, 5, 30, 3.0, 4, arc_deg=40.0), 0),
    "small_huber": (lambda: syn.make_scene("small_huber", 103, 6, 48, 3.5, 4, arc_deg=40.0), 1),
    "k1_quarter": (lambda: syn.config("K1", scale=0.25), 1),
}


def main():
    for name, (build, huber) in CASES.items():
        pb = syn.flatten(build())
        for suffix, prob in (("", pb), ("_g2", fix_second_camera(pb))):
            cam, pt, cost, res = minimise(prob, bool(huber), name + suffix)
            np.savez_compressed(
                os.path.join(HERE, f"scipy_{name}{suffix}.npz"),
                cam=prob.cam, cam_fixed=prob.cam_fixed, fixed_extr=prob.fixed_extr, K=prob.K, pt=prob.pt, obs_cam=prob.obs_cam,
                obs_pt=prob.obs_pt, obs_uv=prob.obs_uv, use_huber=np.int32(huber), opt_cam=cam, opt_pt=pt,
                opt_cost=np.float64(cost), optimality=np.float64(res.optimality), nfev=np.int32(res.nfev),
                gauge_fixed=np.int32(suffix == "_g2"))


if __name__ == "__main__":
    main()
