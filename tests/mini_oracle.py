"""Independent numpy mini-oracle for small problems (<= a few hundred parameters).

Purpose: cross-check the C++ oracle's linear algebra (Schur elimination, back-substitution, Jacobians) with a
completely different route -- DENSE Jacobian by complex-step differentiation of a numpy transcription of the
functors (Optimizer.h:54-76, 96-107), dense normal equations solved with numpy -- under the same trust-region
bookkeeping.  It shares no code with oracle/ba_oracle.cpp.
"""
from __future__ import annotations

import numpy as np

EPS = np.finfo(np.float64).eps


def rodrigues(w):
    """ceres::AngleAxisToRotationMatrix, complex-safe."""
    th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2]
    if np.real(th2) > EPS:
        th = np.sqrt(th2)
        wx, wy, wz = w[0] / th, w[1] / th, w[2] / th
        c, s = np.cos(th), np.sin(th)
        return np.array([
            [c + wx * wx * (1 - c), wx * wy * (1 - c) - wz * s, wy * s + wx * wz * (1 - c)],
            [wz * s + wx * wy * (1 - c), c + wy * wy * (1 - c), -wx * s + wy * wz * (1 - c)],
            [-wy * s + wx * wz * (1 - c), wx * s + wy * wz * (1 - c), c + wz * wz * (1 - c)]])
    one = w[0] * 0 + 1.0
    return np.array([[one, -w[2], w[1]], [w[2], one, -w[0]], [-w[1], w[0], one]])


class MiniBA:
    def __init__(self, pb, use_huber=True, delta=np.sqrt(5.991)):
        self.pb = pb
        self.use_huber = use_huber
        self.a = float(delta)
        Nc, Np, No = pb.shape
        cnt_p = np.bincount(pb.obs_pt, minlength=Np)
        cnt_c = np.bincount(pb.obs_cam, minlength=Nc)
        off = 0
        self.pt_off = np.full(Np, -1)
        self.cam_off = np.full(Nc, -1)
        for p in range(Np):
            if cnt_p[p] > 0:
                self.pt_off[p] = off
                off += 3
        for c in range(Nc):
            if cnt_c[c] > 0 and not pb.cam_fixed[c]:
                self.cam_off[c] = off
                off += 6
        self.n = off

    def pack(self, cam, pt):
        x = np.zeros(self.n)
        for p, o in enumerate(self.pt_off):
            if o >= 0:
                x[o:o + 3] = pt[p]
        for c, o in enumerate(self.cam_off):
            if o >= 0:
                x[o:o + 6] = cam[c]
        return x

    def unpack(self, x):
        cam, pt = self.pb.cam.copy(), self.pb.pt.copy()
        for p, o in enumerate(self.pt_off):
            if o >= 0:
                pt[p] = x[o:o + 3]
        for c, o in enumerate(self.cam_off):
            if o >= 0:
                cam[c] = x[o:o + 6]
        return cam, pt

    def raw_residuals(self, x):
        pb = self.pb
        No = pb.obs_cam.shape[0]
        r = np.zeros(2 * No, dtype=x.dtype)
        for i in range(No):
            c, p = pb.obs_cam[i], pb.obs_pt[i]
            X = x[self.pt_off[p]:self.pt_off[p] + 3]
            K = pb.K[c].astype(np.float64).reshape(3, 3)
            if pb.cam_fixed[c]:
                E = pb.fixed_extr[c].astype(np.float64).reshape(4, 4)
                ph = E[:, :3] @ X + E[:, 3]
                pc = ph[:3] / ph[3]
            else:
                cm = x[self.cam_off[c]:self.cam_off[c] + 6]
                pc = rodrigues(cm[:3]) @ X + cm[3:]
            q = K @ pc
            r[2 * i] = q[0] / q[2] - float(pb.obs_uv[i, 0])
            r[2 * i + 1] = q[1] / q[2] - float(pb.obs_uv[i, 1])
        return r

    def jacobian(self, x):
        h = 1e-30
        J = np.zeros((2 * self.pb.obs_cam.shape[0], self.n))
        for j in range(self.n):
            xc = x.astype(np.complex128)
            xc[j] += 1j * h
            J[:, j] = np.imag(self.raw_residuals(xc)) / h
        return J

    def robust(self, r, J=None):
        """Corrector with alpha = 0 (Huber): returns cost, corrected residuals, corrected Jacobian."""
        r = r.copy()
        s = r[0::2] ** 2 + r[1::2] ** 2
        if not self.use_huber:
            return 0.5 * s.sum(), r, J
        b = self.a * self.a
        out = s > b
        rt = np.sqrt(np.where(out, s, 1.0))
        rho0 = np.where(out, 2 * self.a * rt - b, s)
        rho1 = np.where(out, np.maximum(np.finfo(float).tiny, self.a / rt), 1.0)
        sr = np.repeat(np.sqrt(rho1), 2)
        r *= sr
        if J is not None:
            J = J * sr[:, None]
        return 0.5 * rho0.sum(), r, J

    def cost(self, x):
        return self.robust(self.raw_residuals(x))[0]


def solve(pb, max_num_iterations=50, use_huber=True, function_tolerance=1e-6, gradient_tolerance=1e-10,
          parameter_tolerance=1e-8, radius=1e4, min_relative_decrease=1e-3):
    m = MiniBA(pb, use_huber)
    x = m.pack(pb.cam, pb.pt)
    n = m.n
    its = []

    def linearise(x, scale):
        cost, r, J = m.robust(m.raw_residuals(x), m.jacobian(x))
        g = J.T @ r
        if scale is None:
            scale = 1.0 / (1.0 + np.sqrt((J * J).sum(0)))
        J = J * scale[None]
        d = x - (x + (-g))
        return cost, r, J, scale, np.abs(d).max(), np.linalg.norm(d)

    cost, r, J, scale, gmax, gnorm = linearise(x, None)
    it = dict(iteration=0, cost=cost, cost_change=0.0, gradient_max_norm=gmax, gradient_norm=gnorm,
              step_norm=0.0, relative_decrease=0.0, trust_region_radius=radius, step_is_valid=1,
              step_is_successful=1)
    x_norm = np.linalg.norm(x)
    decrease_factor, reuse, diag = 2.0, False, None
    best_x, min_cost = x.copy(), np.inf
    invalid = 0
    term = None
    while True:
        if it["step_is_successful"] and cost < min_cost:
            min_cost, best_x = cost, x.copy()
        it["trust_region_radius"] = radius
        its.append(it)
        if it["iteration"] >= max_num_iterations:
            term = "NO_CONVERGENCE"; break
        if it["step_is_successful"] and it["gradient_max_norm"] <= gradient_tolerance:
            term = "CONVERGENCE:gradient"; break
        if radius <= 1e-32:
            term = "CONVERGENCE:radius"; break
        prev = it
        it = dict(iteration=prev["iteration"] + 1, cost=0.0, cost_change=0.0, gradient_max_norm=0.0,
                  gradient_norm=0.0, step_norm=0.0, relative_decrease=0.0, trust_region_radius=0.0,
                  step_is_valid=0, step_is_successful=0)
        if not reuse:
            diag = np.clip((J * J).sum(0), 1e-6, 1e32)
        D = np.sqrt(diag / radius)
        A = J.T @ J + np.diag(D * D)
        try:
            L = np.linalg.cholesky(A)
            y = np.linalg.solve(L.T, np.linalg.solve(L, J.T @ r))
            ok = np.all(np.isfinite(y))
        except np.linalg.LinAlgError:
            ok = False
        reuse = True
        mcc = 0.0
        if ok:
            step = -y
            mr = J @ step
            mcc = -mr @ (r + mr / 2.0)
            it["step_is_valid"] = int(mcc > 0)
        if not it["step_is_valid"]:
            invalid += 1
            if invalid >= 5:
                term = "FAILURE"; break
            radius /= decrease_factor; decrease_factor *= 2.0
            it.update(cost=cost, gradient_max_norm=prev["gradient_max_norm"], gradient_norm=prev["gradient_norm"])
            continue
        invalid = 0
        delta = step * scale
        cand = x + delta
        cand_cost = m.cost(cand)
        it["step_norm"] = np.linalg.norm(x - cand)
        if it["step_norm"] <= parameter_tolerance * (x_norm + parameter_tolerance):
            term = "CONVERGENCE:parameter"; break
        it["cost_change"] = cost - cand_cost
        if abs(it["cost_change"]) <= function_tolerance * cost:
            term = "CONVERGENCE:function"; break
        it["relative_decrease"] = (cost - cand_cost) / mcc
        if it["relative_decrease"] > min_relative_decrease:
            x = cand
            x_norm = np.linalg.norm(x)
            cost, r, J, scale, gmax, gnorm = linearise(x, scale)
            it.update(cost=cost, gradient_max_norm=gmax, gradient_norm=gnorm, step_is_successful=1)
            radius = min(1e16, radius / max(1.0 / 3.0, 1.0 - (2.0 * it["relative_decrease"] - 1.0) ** 3))
            decrease_factor, reuse = 2.0, False
        else:
            it.update(cost=cand_cost, gradient_max_norm=prev["gradient_max_norm"], gradient_norm=prev["gradient_norm"])
            radius /= decrease_factor; decrease_factor *= 2.0
    cam, pt = m.unpack(best_x)
    return cam, pt, term, its
