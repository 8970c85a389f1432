"""Cross-checks of the C++ oracle against the independent numpy mini-oracle (tests/mini_oracle.py: complex-step
dense Jacobian + dense normal equations -- shares no code with oracle/ba_oracle.cpp), plus internal consistency of
the oracle's own alternative routes (PCG vs dense, shard sums, pose-only batch vs the general solver).  CPU only."""
import dataclasses

import numpy as np
import pytest

import mini_oracle
from bundleadjustment_b200 import synthetic as syn


def _small(seed, n_cam=5, n_pt=40):
    return syn.flatten(syn.make_scene("t", seed, n_cam, n_pt, 3.0, 4, arc_deg=40.0))


@pytest.mark.parametrize("seed,huber", [(1, True), (2, False), (3, True)])
def test_trajectory_matches_mini_oracle(orc, seed, huber):
    pb = _small(seed)
    cam, pt, sm, its = orc.solve(pb, orc.default_options(max_num_iterations=25, use_huber=int(huber)))
    mcam, mpt, term, mits = mini_oracle.solve(pb, 25, use_huber=huber)
    assert len(its) == len(mits)
    for a, b in zip(its, mits):
        assert a["step_is_successful"] == b["step_is_successful"]
        assert abs(a["cost"] - b["cost"]) <= 1e-7 * b["cost"]
        assert abs(a["trust_region_radius"] - b["trust_region_radius"]) <= 1e-5 * b["trust_region_radius"]
        assert abs(a["gradient_max_norm"] - b["gradient_max_norm"]) <= 1e-5 * max(b["gradient_max_norm"], 1e-6)
    assert (sm.termination_type == orc.CONVERGENCE) == term.startswith("CONVERGENCE")
    assert np.abs(cam - mcam).max() <= 1e-6 and np.abs(pt - mpt).max() <= 1e-6


def test_jacobian_matches_complex_step(orc):
    pb = _small(4, 4, 25)
    ev = orc.evaluate(pb, orc.default_options(use_huber=0))
    m = mini_oracle.MiniBA(pb, use_huber=False)
    x = m.pack(pb.cam, pb.pt)
    J = m.jacobian(x)
    r = m.raw_residuals(x)
    assert np.abs(r - ev["residuals"].reshape(-1)).max() <= 1e-11
    for i in range(pb.obs_cam.shape[0]):
        c, p = pb.obs_cam[i], pb.obs_pt[i]
        Jp = J[2 * i:2 * i + 2, m.pt_off[p]:m.pt_off[p] + 3]
        assert np.abs(Jp - ev["jac_pt"][i]).max() <= 1e-10 * np.abs(Jp).max()
        if m.cam_off[c] >= 0:
            Jc = J[2 * i:2 * i + 2, m.cam_off[c]:m.cam_off[c] + 6]
            assert np.abs(Jc - ev["jac_cam"][i]).max() <= 1e-10 * np.abs(Jc).max()
        else:
            assert not ev["jac_cam"][i].any()


def test_schur_system_equals_dense_elimination(orc):
    """S and rhs from the oracle's per-landmark elimination == the Schur complement of the dense damped normal
    equations built from its own materialised Jacobian."""
    pb = _small(5, 5, 30)
    Nc, Np, No = pb.shape
    radius = 123.0
    ev = orc.evaluate(pb)
    J = np.zeros((2 * No, 3 * Np + 6 * Nc))
    for i in range(No):
        J[2 * i:2 * i + 2, 3 * pb.obs_pt[i]:3 * pb.obs_pt[i] + 3] = ev["jac_pt"][i]
        J[2 * i:2 * i + 2, 3 * Np + 6 * pb.obs_cam[i]:3 * Np + 6 * pb.obs_cam[i] + 6] = ev["jac_cam"][i]
    r = ev["residuals"].reshape(-1)
    live = (J * J).sum(0) > 0
    scale = np.where(live, 1.0 / (1.0 + np.sqrt((J * J).sum(0))), 1.0)
    Js = J * scale
    H = Js.T @ Js
    d = np.clip(np.diag(H), 1e-6, 1e32) / radius
    H = H + np.diag(np.where(live, d, 0.0))
    g = Js.T @ r
    P, Cc = slice(0, 3 * Np), slice(3 * Np, None)
    lp = live[P]
    Hpp = H[P, P][np.ix_(lp, lp)]
    Hcp = H[Cc, P][:, lp]
    S = H[Cc, Cc] - Hcp @ np.linalg.solve(Hpp, Hcp.T)
    rhs = g[Cc] - Hcp @ np.linalg.solve(Hpp, g[P][lp])
    So, rhso, cost = orc.schur_system(pb, radius, 0, Np)
    Su = np.triu(So) + np.triu(So, 1).T
    lc = live[Cc]
    assert np.abs(Su[np.ix_(lc, lc)] - S[np.ix_(lc, lc)]).max() <= 1e-9 * np.abs(S).max()
    assert np.abs(rhso[lc] - rhs[lc]).max() <= 1e-9 * np.abs(rhs).max()
    assert abs(cost - ev["cost"]) <= 1e-13 * cost


def test_schur_shards_sum_to_whole(orc):
    """Landmark sharding: per-shard reduced systems add up to the whole; the replicated camera terms
    (U diagonal blocks, D_c^2, b_c) are contributed by every shard's own observations only."""
    pb = _small(6, 6, 60)
    Np = pb.pt.shape[0]
    S, rhs, cost = orc.schur_system(pb, 1e4, 0, Np)
    cut = [0, 17, 41, Np]
    parts = [orc.schur_system(pb, 1e4, a, b) for a, b in zip(cut, cut[1:])]
    assert abs(sum(p[2] for p in parts) - cost) <= 1e-12 * cost
    # off-diagonal blocks and rhs landmark terms are plain sums; the LM diagonal is added once (by the full system)
    Ssum = sum(p[0] for p in parts)
    off = ~np.kron(np.eye(pb.cam.shape[0], dtype=bool), np.ones((6, 6), dtype=bool))
    assert np.abs(Ssum[off] - S[off]).max() <= 1e-10 * np.abs(S).max()


def test_oracle_pcg_tight_equals_dense(orc):
    pb = syn.flatten(syn.config("K1", scale=0.1))
    dense = orc.solve(pb, orc.default_options(max_num_iterations=15))
    pcg = orc.solve(pb, orc.default_options(max_num_iterations=15, linear_solver=orc.ITERATIVE_SCHUR, eta=1e-14,
                                            cg_r_tolerance=1e-13, max_linear_solver_iterations=3000))
    assert len(dense[3]) == len(pcg[3])
    for a, b in zip(dense[3], pcg[3]):
        assert abs(a["cost"] - b["cost"]) <= 1e-7 * a["cost"]
    assert abs(dense[2].final_cost - pcg[2].final_cost) <= 1e-8 * dense[2].final_cost
    # Ceres' own stopping rule (eta = 0.1) is inexact but still converges to the same minimum
    loose = orc.solve(pb, orc.default_options(max_num_iterations=60, linear_solver=orc.ITERATIVE_SCHUR))
    assert loose[2].termination_type == orc.CONVERGENCE
    assert abs(loose[2].final_cost - dense[2].final_cost) <= 1e-3 * dense[2].final_cost
    assert all(i["linear_solver_iterations"] >= 1 for i in loose[3][1:])


def test_pose_only_batch_equals_general_solver(orc):
    """MotionOnly BA (one 6-dof block, landmarks baked in as float32): iteration-0 cost by hand and the first LM
    step from the 6x6 damped normal equations (J^T J + D^2) y = J^T r assembled in numpy."""
    batch = syn.make_pose_batch(seed=3, n_prob=4, n_pts=60)
    cam, sms, iters = orc.solve_pose_only_batch(batch, orc.default_options(max_num_iterations=20))
    for b in range(4):
        a, e = batch.prob_ptr[b], batch.prob_ptr[b + 1]
        x = batch.cam[b].copy()
        # iteration-0 cost by hand
        rs = [orc.eval_pose_only(x, batch.Xh[i], batch.uv[i], batch.K[b])[0] for i in range(a, e)]
        s = np.array([(r ** 2).sum() for r in rs])
        A2 = 5.991
        rho = np.where(s > A2, 2 * np.sqrt(A2 * s) - A2, s)
        assert abs(iters[b][0]["cost"] - 0.5 * rho.sum()) <= 1e-12 * iters[b][0]["cost"]
        assert sms[b].termination_type == orc.CONVERGENCE
        assert sms[b].final_cost < 0.5 * sms[b].initial_cost
        ev = [orc.eval_pose_only(x, batch.Xh[i], batch.uv[i], batch.K[b]) for i in range(a, e)]
        sr = np.sqrt(np.where(s > A2, np.sqrt(A2 / np.where(s > A2, s, 1.0)), 1.0))
        J = np.concatenate([ev[k][1] * sr[k] for k in range(e - a)])
        r = np.concatenate([ev[k][0] * sr[k] for k in range(e - a)])
        scale = 1.0 / (1.0 + np.sqrt((J * J).sum(0)))
        Js = J * scale
        D2 = np.clip((Js * Js).sum(0), 1e-6, 1e32) / 1e4
        y = np.linalg.solve(Js.T @ Js + np.diag(D2), Js.T @ r)
        xc = x - y * scale
        rs1 = [orc.eval_pose_only(xc, batch.Xh[i], batch.uv[i], batch.K[b])[0] for i in range(a, e)]
        s1 = np.array([(q ** 2).sum() for q in rs1])
        c1 = 0.5 * np.where(s1 > A2, 2 * np.sqrt(A2 * s1) - A2, s1).sum()
        assert iters[b][1]["step_is_successful"] == 1
        assert abs(iters[b][1]["cost"] - c1) <= 1e-9 * c1
        assert np.abs(cam[b] - batch.gt_cam[b]).max() < 0.02
    # a masked-out observation set: all masked -> nothing to do, parameters untouched
    mask = np.zeros(batch.uv.shape[0], np.uint8)
    cam2, sms2, _ = orc.solve_pose_only_batch(batch, orc.default_options(max_num_iterations=20), mask=mask)
    assert np.array_equal(cam2, batch.cam)


def test_prune_frame_semantics(orc):
    """pruneCorrespondences (Optimizer.cpp:6-79): cheirality, distance gate, NON-squared pixel distance over
    1.2^octave against 5.991, consider_outlier switch."""
    pose = np.eye(4, dtype=np.float32)
    K = np.array([[500, 0, 320], [0, 500, 240], [0, 0, 1]], np.float32)
    X = np.array([[0, 0, 5], [0, 0, -5], [0, 0, 5], [0, 0, 5], [0.06, 0, 5], [0.06, 0, 5], [0, 0, 5], [0, 0, 5]], np.float32)
    kp = np.tile(np.array([[320, 240]], np.float32), (8, 1))
    octave = np.array([0, 0, 0, 0, 0, 1, 0, 0], np.int32)
    mn = np.array([1, 1, 6, 1, 1, 1, 1, 1], np.float32)
    mx = np.array([9, 9, 9, 4, 9, 9, 9, 9], np.float32)
    has = np.array([1, 1, 1, 1, 1, 1, 0, 1], np.uint8)
    out = np.array([0, 0, 0, 0, 0, 0, 1, 1], np.uint8)
    got = orc.prune_frame(pose, K, kp, octave, X, mn, mx, has, out, consider_outlier=True)
    # 0 inlier; 1 behind the camera; 2 nearer than minDist; 3 beyond maxDist; 4: 6 px off at octave 0 -> 6 > 5.991
    # outlier; 5: same 6 px at octave 1 -> 6 / 1.2 = 5 inlier; 6 has no landmark -> untouched; 7 outlier -> inlier again
    assert got.tolist() == [0, 1, 1, 1, 1, 0, 1, 0]
    got = orc.prune_frame(pose, K, kp, octave, X, mn, mx, has, out, consider_outlier=False)
    assert got.tolist() == [0, 1, 1, 1, 1, 0, 1, 1]          # 7 stays an outlier: outliers are not revisited


def test_flatten_write_back_round_trip(orc):
    """Optimizer.cpp:296-299 / 262-267: pose -> (angle-axis, t) -> pose is the identity up to float32 rounding."""
    sc = syn.config("K1", scale=0.2)
    for f in range(sc.pose_f32.shape[0]):
        cam, extr = orc.pose_to_cam(sc.pose_f32[f])
        assert np.abs(extr - np.linalg.inv(sc.pose_f32[f].astype(np.float64))).max() < 2e-6
        pose = orc.cam_to_pose(cam)
        assert np.abs(pose - sc.pose_f32[f]).max() < 5e-6
    pb = syn.flatten(sc)
    cam0, _ = orc.pose_to_cam(sc.pose_f32[3])
    assert np.abs(cam0 - pb.cam[3]).max() < 1e-6
