"""Known-answer tests that pin the CPU oracle (the reference has NO tests / golden vectors for the BA
path, so these are the hand-derived KATs the survey lists).  CPU only.

(1) Rodrigues incl. the theta^2 <= eps branch and the column-major layout (ceres/rotation.h semantics).
(2) The functors on the reference's own debug geometry SfMHelper::setUpDebugCam (SfMHelper.cpp:155-201): f = 100,
    c = (320, 240), rotation 45 deg about x, t = (0, 0, 1); pixel values computed by hand below.
(3) Analytic Jacobian == dual-number Jacobian (autodiff stand-in) and ~ central differences.
(4) Huber just below / above 5.991.
(5) A one-camera / one-landmark LM step by hand.
(6) Termination edge cases.
"""
import dataclasses

import numpy as np
import pytest

from bundleadjustment_b200 import synthetic as syn

A = np.sqrt(5.991)


# ---------------------------------------------------------------------------------------------------- (1)
def test_rodrigues_known_answers(orc):
    assert np.array_equal(orc.angle_axis_to_rotation([0, 0, 0]), np.eye(3))
    R = orc.angle_axis_to_rotation([np.pi / 2, 0, 0])
    want = np.array([[1, 0, 0], [0, 0, -1], [0, 1, 0]], float)       # rotation about x by +90 deg
    assert np.abs(R - want).max() < 1e-15
    # column-major layout: the binding transposes; R[1,2] (row 1, col 2) must be -sin
    assert R[1, 2] < -0.99 and R[2, 1] > 0.99
    # small-angle branch: theta^2 <= DBL_EPSILON -> R = I + [w]x exactly (first order)
    w = np.array([1e-9, -2e-9, 3e-9])
    R = orc.angle_axis_to_rotation(w)
    want = np.array([[1, -w[2], w[1]], [w[2], 1, -w[0]], [-w[1], w[0], 1]])
    assert np.array_equal(R, want)
    # just above the branch point the closed form is used and is orthonormal
    w = np.array([3e-8, 0, 0])
    R = orc.angle_axis_to_rotation(w)
    assert np.abs(R @ R.T - np.eye(3)).max() < 1e-15


def test_rotation_round_trip(orc):
    rng = np.random.default_rng(0)
    for _ in range(200):
        w = rng.normal(size=3) * rng.choice([1e-10, 1e-3, 0.5, 2.5])
        R = orc.angle_axis_to_rotation(w)
        w2 = orc.rotation_to_angle_axis(R)
        assert np.abs(orc.angle_axis_to_rotation(w2) - R).max() < 1e-12
    # pi rotation: axis sign is ambiguous, the matrix is not
    R = orc.angle_axis_to_rotation([0, np.pi, 0])
    assert np.abs(orc.angle_axis_to_rotation(orc.rotation_to_angle_axis(R)) - R).max() < 1e-12


# ---------------------------------------------------------------------------------------------------- (2)
def _debug_cam():
    K = np.array([[100, 0, 320], [0, 100, 240], [0, 0, 1]], np.float32)
    cam = np.array([np.pi / 4, 0, 0, 0, 0, 1.0])
    return K, cam


def test_functor_debug_cam_pixels(orc):
    """setUpDebugCam geometry.  X = (0,0,10): p = R X + t = (0, -10 s, 10 c + 1), s = c = sqrt(1/2)
    -> u = 320, v = 240 - 100 * (10 s) / (10 c + 1) ;  X = (1,0,10): u = 320 + 100 / (10 c + 1)."""
    K, cam = _debug_cam()
    s = c = np.sqrt(0.5)
    z = 10 * c + 1
    for X, (u, v) in [((0, 0, 10), (320.0, 240 - 100 * 10 * s / z)), ((1, 0, 10), (320 + 100 / z, 240 - 100 * 10 * s / z)),
                      ((0, 1, 10), (320.0, 240 + 100 * (c - 10 * s) / (s + 10 * c + 1)))]:
        uv = np.array([u, v], np.float32)
        r, Jc, Jp = orc.eval_cam_point(cam, X, uv, K)
        # residual = predicted - float32(observed): zero up to the float32 rounding of the observation
        assert np.abs(r - (np.array([u, v]) - uv.astype(np.float64))).max() < 1e-12
        # first camera of the pair (identity): u = 320 + 100 x / z, v = 240 + 100 y / z
        r0, _, _ = orc.eval_cam_point(np.zeros(6), X, np.array([0, 0], np.float32), K)
        assert np.allclose(r0, [320 + 100 * X[0] / X[2], 240 + 100 * X[1] / X[2]], rtol=0, atol=1e-12)


def test_point_only_and_pose_only_agree_with_cam_point(orc):
    """R2 with the extrinsics of a camera baked in (float32 4x4) and R3 with the landmark baked in (float32
    homogeneous) are the same projection as R1 when the baked values are exactly representable."""
    K, _ = _debug_cam()
    cam = np.array([0.0, 0.0, 0.0, 0.25, -0.5, 1.0])                  # identity rotation, dyadic translation
    E = np.eye(4, dtype=np.float32); E[:3, 3] = cam[3:]
    X = np.array([0.5, -0.25, 4.0])
    uv = np.array([330.5, 225.25], np.float32)
    r1, Jc1, Jp1 = orc.eval_cam_point(cam, X, uv, K)
    r2, Jp2 = orc.eval_point_only(X, uv, E, K)
    r3, Jc3 = orc.eval_pose_only(cam, np.array([*X, 1.0], np.float32), uv, K)
    assert np.abs(r1 - r2).max() < 1e-13 and np.abs(Jp1 - Jp2).max() < 1e-12
    assert np.abs(r1 - r3).max() < 1e-13 and np.abs(Jc1 - Jc3).max() < 1e-12
    # R2 divides by the 4th homogeneous coordinate (Optimizer.h:100 hnormalized): a last row != (0,0,0,1) matters
    E2 = E.copy(); E2[3, 3] = 2.0
    r4, _ = orc.eval_point_only(X, uv, E2, K)
    p = (E[:3, :3].astype(float) @ X + E[:3, 3]) / 2.0
    want = np.array([100 * p[0] / p[2] + 320, 100 * p[1] / p[2] + 240]) - uv
    assert np.abs(r4 - want).max() < 1e-12


# ---------------------------------------------------------------------------------------------------- (3)
@pytest.mark.parametrize("rot_scale", [0.0, 1e-9, 1e-4, 0.3, 2.0])
def test_analytic_jacobian_equals_jets_and_differences(orc, rot_scale):
    rng = np.random.default_rng(int(rot_scale * 1e6) + 1)
    K = np.array([[525, 0, 319.5], [0, 525, 239.5], [0, 0, 1]], np.float32)
    for _ in range(20):
        cam = np.concatenate([rng.normal(size=3) * rot_scale, rng.normal(size=3) * 0.3])
        # landmark placed in front of THIS camera: X = R^T (Xc - t)
        Xc = np.array([rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(2, 6)])
        X = orc.angle_axis_to_rotation(cam[:3]).T @ (Xc - cam[3:])
        uv = rng.uniform(0, 480, size=2).astype(np.float32)
        r, Jc, Jp = orc.eval_cam_point(cam, X, uv, K)
        rj, Jcj, Jpj = orc.eval_cam_point(cam, X, uv, K, jets=True)
        assert np.abs(r - rj).max() <= 1e-13 * max(np.abs(r).max(), 1.0)
        assert np.abs(Jc - Jcj).max() <= 1e-13 * max(np.abs(Jcj).max(), 1.0)
        assert np.abs(Jp - Jpj).max() <= 1e-13 * max(np.abs(Jpj).max(), 1.0)
        if rot_scale in (0.0, 1e-9):
            continue        # the small-angle branch differentiates the first-order formula; differences see sin/cos
        h = 1e-6
        for j in range(6):
            d = np.zeros(6); d[j] = h
            fd = (orc.eval_cam_point(cam + d, X, uv, K)[0] - orc.eval_cam_point(cam - d, X, uv, K)[0]) / (2 * h)
            assert np.abs(fd - Jc[:, j]).max() <= 1e-6 * max(np.abs(Jc).max(), 1.0)
        for j in range(3):
            d = np.zeros(3); d[j] = h
            fd = (orc.eval_cam_point(cam, X + d, uv, K)[0] - orc.eval_cam_point(cam, X - d, uv, K)[0]) / (2 * h)
            assert np.abs(fd - Jp[:, j]).max() <= 1e-6 * max(np.abs(Jp).max(), 1.0)
        # the two one-block functors
        E = np.eye(4, dtype=np.float32)
        E[:3, :3] = orc.angle_axis_to_rotation(cam[:3]).astype(np.float32); E[:3, 3] = cam[3:].astype(np.float32)
        E[3] = [1e-4, -2e-4, 3e-4, 1.0]                  # "approximately 0 0 0 1" (float inverse residue)
        r2, Jp2 = orc.eval_point_only(X, uv, E, K)
        r2j, Jp2j = orc.eval_point_only(X, uv, E, K, jets=True)
        assert np.abs(Jp2 - Jp2j).max() <= 1e-13 * np.abs(Jp2j).max()
        Xh = np.array([*X, 1.0], np.float32)
        r3, Jc3 = orc.eval_pose_only(cam, Xh, uv, K)
        r3j, Jc3j = orc.eval_pose_only(cam, Xh, uv, K, jets=True)
        assert np.abs(Jc3 - Jc3j).max() <= 1e-13 * np.abs(Jc3j).max()


# ---------------------------------------------------------------------------------------------------- (4)
def test_huber_known_answers(orc):
    b = A * A
    assert abs(b - 5.991) < 1e-15
    assert np.array_equal(orc.huber(0.0, A), [0.0, 1.0, 0.0])
    assert np.array_equal(orc.huber(5.99, A), [5.99, 1.0, 0.0])
    assert np.array_equal(orc.huber(b, A), [b, 1.0, 0.0])                       # s <= b is the inlier region
    s = 5.992
    r = np.sqrt(s)
    rho = orc.huber(s, A)
    assert np.allclose(rho, [2 * A * r - b, A / r, -(A / r) / (2 * s)], rtol=1e-15, atol=0)
    s = 100.0
    assert np.allclose(orc.huber(s, A), [2 * A * 10 - b, A / 10, -A / 10 / 200], rtol=1e-15, atol=0)


def test_robustified_cost_and_corrector(orc):
    """cost = 1/2 sum rho(s); corrected residual = sqrt(rho') r, corrected J = sqrt(rho') J (alpha = 0 branch)."""
    pb = syn.flatten(syn.make_scene("t", 3, 5, 60, 3.0, 4, arc_deg=40.0))
    ev = orc.evaluate(pb)
    raw = orc.evaluate(pb, orc.default_options(use_huber=0))
    s = (raw["residuals"] ** 2).sum(1)
    out = s > A * A
    assert out.any() and (~out).any()
    rho0 = np.where(out, 2 * A * np.sqrt(s) - A * A, s)
    assert abs(ev["cost"] - 0.5 * rho0.sum()) <= 1e-13 * ev["cost"]
    sr = np.sqrt(np.where(out, A / np.sqrt(s), 1.0))
    assert np.abs(ev["residuals"] - raw["residuals"] * sr[:, None]).max() < 1e-13
    assert np.abs(ev["jac_pt"] - raw["jac_pt"] * sr[:, None, None]).max() <= 1e-13 * np.abs(raw["jac_pt"]).max()
    # gradient = J^T r over the corrected blocks
    g = np.zeros_like(pb.pt)
    np.add.at(g, pb.obs_pt, np.einsum('nij,ni->nj', ev["jac_pt"], ev["residuals"]))
    assert np.abs(g - ev["grad_pt"]).max() <= 1e-12 * np.abs(g).max()


# ---------------------------------------------------------------------------------------------------- (5)
def test_single_lm_step_by_hand(orc):
    """Hand-assembled first LM step: Jacobi scale from iteration 0, D = sqrt(clamp(diag) / radius) on the SCALED
    Jacobian, dense normal equations, step = -y, candidate = x + step * scale -- compared with orc_solve's
    iteration 1 (cost, step norm, rho, radius)."""
    pb = syn.flatten(syn.make_scene("t", 21, 3, 4, 3.0, 2, arc_deg=30.0))
    ev = orc.evaluate(pb)
    Nc, Np, No = pb.shape
    free = [c for c in range(Nc) if not pb.cam_fixed[c] and (pb.obs_cam == c).any()]
    act = [p for p in range(Np) if (pb.obs_pt == p).any()]
    n = 0
    offs = {}
    for p in act:
        offs[("p", p)] = n; n += 3
    for c in free:
        offs[("c", c)] = n; n += 6
    J = np.zeros((2 * No, n))
    r = ev["residuals"].reshape(-1)
    for i in range(No):
        c, p = int(pb.obs_cam[i]), int(pb.obs_pt[i])
        J[2 * i:2 * i + 2, offs[("p", p)]:offs[("p", p)] + 3] = ev["jac_pt"][i]
        if ("c", c) in offs:
            J[2 * i:2 * i + 2, offs[("c", c)]:offs[("c", c)] + 6] = ev["jac_cam"][i]
    scale = 1.0 / (1.0 + np.sqrt((J * J).sum(0)))
    Js = J * scale
    radius = 1e4
    D2 = np.clip((Js * Js).sum(0), 1e-6, 1e32) / radius
    y = np.linalg.solve(Js.T @ Js + np.diag(D2), Js.T @ r)
    step = -y
    mcc = -(Js @ step) @ (r + 0.5 * (Js @ step))
    delta = step * scale
    cam, pt = pb.cam.copy(), pb.pt.copy()
    for p in act:
        pt[p] += delta[offs[("p", p)]:offs[("p", p)] + 3]
    for c in free:
        cam[c] += delta[offs[("c", c)]:offs[("c", c)] + 6]
    cand_cost = orc.evaluate(pb, cam=cam, pt=pt)["cost"]
    rho = (ev["cost"] - cand_cost) / mcc
    _, _, sm, its = orc.solve(pb, orc.default_options(max_num_iterations=1))
    assert len(its) == 2 and its[0]["iteration"] == 0 and its[1]["iteration"] == 1
    assert abs(its[0]["cost"] - ev["cost"]) <= 1e-14 * ev["cost"]
    assert its[0]["trust_region_radius"] == 1e4
    g = J.T @ r
    assert abs(its[0]["gradient_max_norm"] - np.abs(g).max()) <= 1e-10 * np.abs(g).max()
    assert mcc > 0 and its[1]["step_is_valid"] == 1
    assert abs(its[1]["relative_decrease"] - rho) <= 1e-7 * abs(rho)
    assert abs(its[1]["step_norm"] - np.linalg.norm(delta)) <= 1e-8 * np.linalg.norm(delta)
    if rho > 1e-3:
        assert its[1]["step_is_successful"] == 1
        assert abs(its[1]["cost"] - cand_cost) <= 1e-9 * cand_cost
        want_radius = min(1e16, radius / max(1.0 / 3.0, 1.0 - (2 * rho - 1) ** 3))
        assert abs(its[1]["trust_region_radius"] - want_radius) <= 1e-6 * want_radius


# ---------------------------------------------------------------------------------------------------- (6)
def test_termination_edge_cases(orc):
    pb = syn.flatten(syn.make_scene("t", 7, 5, 80, 3.0, 4, arc_deg=40.0))
    # max_num_iterations = 0: only the initial evaluation is recorded
    cam, pt, sm, its = orc.solve(pb, orc.default_options(max_num_iterations=0))
    assert len(its) == 1 and sm.termination_type == orc.NO_CONVERGENCE
    assert np.array_equal(cam, pb.cam) and np.array_equal(pt, pb.pt)
    # function tolerance: CONVERGENCE; the last candidate is NOT applied and not recorded, so the recorded costs
    # are monotone and final_cost == min recorded cost
    cam, pt, sm, its = orc.solve(pb, orc.default_options(max_num_iterations=200))
    assert sm.termination_type == orc.CONVERGENCE
    assert b"tolerance" in sm.message.lower() or b"Tolerance" in sm.message
    costs = [i["cost"] for i in its if i["step_is_successful"]]
    assert all(b <= a for a, b in zip(costs, costs[1:]))
    assert sm.final_cost == min(i["cost"] for i in its)
    assert abs(orc.evaluate(pb, cam=cam, pt=pt)["cost"] - sm.final_cost) <= 1e-12 * sm.final_cost
    # a huge function tolerance stops at the first step without taking it
    cam, pt, sm, its = orc.solve(pb, orc.default_options(function_tolerance=10.0))
    assert len(its) == 1 and sm.termination_type == orc.CONVERGENCE
    assert np.array_equal(cam, pb.cam) and np.array_equal(pt, pb.pt)
    # invalid steps: NaN observations poison the step -> FAILURE after max_num_consecutive_invalid_steps, or an
    # evaluation failure at iteration 0; either way parameters are returned untouched
    bad = dataclasses.replace(pb, obs_uv=pb.obs_uv.copy())
    bad.obs_uv[3, 0] = np.nan
    cam, pt, sm, its = orc.solve(bad, orc.default_options(max_num_iterations=20))
    assert sm.termination_type == orc.FAILURE
    assert np.array_equal(cam, pb.cam) and np.array_equal(pt, pb.pt)


def test_radius_shrinks_on_rejected_steps(orc):
    """Rejected step: radius /= decrease_factor, decrease_factor *= 2; an accepted step resets the factor to 2.
    min_relative_decrease = 0.95 forces rejections on an otherwise easy problem (rho of a good step is 0.8-0.99)."""
    pb = syn.flatten(syn.make_scene("t", 8, 4, 30, 3.0, 4, arc_deg=40.0))
    rng = np.random.default_rng(0)
    cam = pb.cam.copy(); cam[1:] += rng.normal(scale=0.6, size=cam[1:].shape)
    pb = dataclasses.replace(pb, cam=cam)
    _, _, sm, its = orc.solve(pb, orc.default_options(max_num_iterations=12, use_huber=0, min_relative_decrease=0.95))
    rej = [k for k in range(1, len(its)) if its[k]["step_is_valid"] and not its[k]["step_is_successful"]]
    assert rej and rej[0] == 1 and 2 in rej, rej
    # its[k].trust_region_radius is the radius AFTER the update of iteration k
    assert its[1]["trust_region_radius"] == its[0]["trust_region_radius"] / 2.0
    assert its[2]["trust_region_radius"] == its[1]["trust_region_radius"] / 4.0
    # a rejected iteration reports the candidate's cost but x (and the next iteration's reference cost) is unchanged
    acc = [k for k in range(1, len(its)) if its[k]["step_is_successful"]]
    assert acc, "no accepted step"
    k = acc[0]
    assert its[k]["cost"] < its[0]["cost"]
    assert abs((its[0]["cost"] - its[k]["cost"]) - its[k]["cost_change"]) <= 1e-9 * its[0]["cost"]
    if k + 1 in rej:
        assert its[k + 1]["trust_region_radius"] == its[k]["trust_region_radius"] / 2.0    # factor was reset
