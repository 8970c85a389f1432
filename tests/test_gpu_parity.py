"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Tolerances (north_star): per-iteration cost trajectory 1e-6 relative, final cost 1e-8 relative, poses / points 1e-6.
Per-entry residual / Jacobian parity is held to 1e-10 (SURVEY 7.2 asks <= 1e-12 relative per entry; entries close
to zero are compared absolutely against the row scale)."""
import numpy as np
import pytest

from bundleadjustment_b200 import synthetic as syn

pytestmark = pytest.mark.gpu


def _traj_check(its, oits, rtol=1e-6):
    assert len(its) == len(oits), (len(its), len(oits), [i["cost"] for i in its], [i["cost"] for i in oits])
    for a, b in zip(its, oits):
        assert a["iteration"] == b["iteration"]
        assert a["step_is_successful"] == b["step_is_successful"], (a, b)
        assert abs(a["cost"] - b["cost"]) <= rtol * abs(b["cost"]), (a, b)
        assert abs(a["trust_region_radius"] - b["trust_region_radius"]) <= 1e-5 * b["trust_region_radius"], (a, b)


@pytest.mark.parametrize("name", ["tiny_huber", "tiny_l2", "small_huber"])
def test_golden_fixtures(bab, name):
    """The CUDA path against the committed golden vectors (tests/golden/make_golden.py)."""
    import golden_util as G
    pb, z, gits = G.load(name)
    s = bab.Solver(bab.default_options(max_num_iterations=int(z["max_num_iterations"]), use_huber=int(z["use_huber"])))
    s.set_problem(pb)
    ev = s.evaluate()
    sm = s.solve()
    G.check_against_golden(z, gits, ev, s.iterations(), s.cameras(), s.points(), sm.final_cost, sm.termination_type)
    s.close()


@pytest.fixture(scope="module")
def k1_small():
    return syn.flatten(syn.config("K1", scale=0.25))


def test_evaluate_matches_oracle(bab, orc, k1_small):
    pb = k1_small
    s = bab.Solver(bab.default_options())
    s.set_problem(pb)
    ev = s.evaluate()
    ref = orc.evaluate(pb)
    assert abs(ev["cost"] - ref["cost"]) <= 1e-13 * ref["cost"]
    for k in ("residuals", "jac_cam", "jac_pt"):
        scale = np.abs(ref[k]).max()
        assert np.abs(ev[k] - ref[k]).max() <= 1e-11 * scale, k
    for k in ("grad_cam", "grad_pt"):
        scale = np.abs(ref[k]).max()
        assert np.abs(ev[k] - ref[k]).max() <= 1e-10 * scale, k
    s.close()


def test_evaluate_shuffled_observation_order(bab, orc, k1_small):
    """Outputs come back in the CALLER's observation order whatever the internal landmark-major layout is."""
    pb = k1_small
    rng = np.random.default_rng(0)
    perm = rng.permutation(pb.obs_cam.shape[0])
    import dataclasses
    pb2 = dataclasses.replace(pb, obs_cam=pb.obs_cam[perm].copy(), obs_pt=pb.obs_pt[perm].copy(), obs_uv=pb.obs_uv[perm].copy())
    s = bab.Solver(bab.default_options())
    s.set_problem(pb2)
    ev = s.evaluate(gradient=False)
    ref = orc.evaluate(pb2)
    assert np.abs(ev["residuals"] - ref["residuals"]).max() <= 1e-10
    assert np.abs(ev["jac_cam"] - ref["jac_cam"]).max() <= 1e-11 * np.abs(ref["jac_cam"]).max()
    s.close()


def test_no_huber_and_fixed_only_landmarks(bab, orc):
    sc = syn.make_scene("t", 11, 6, 200, 3.0, 4, arc_deg=40.0)
    pb = syn.flatten(sc)
    opt = bab.default_options(use_huber=0, max_num_iterations=15)
    s = bab.Solver(opt)
    s.set_problem(pb)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb, orc.default_options(use_huber=0, max_num_iterations=15))
    _traj_check(s.iterations(), oits)
    assert sm.termination_type == osm.termination_type
    assert np.abs(s.cameras() - cam).max() <= 1e-6
    assert np.abs(s.points() - pt).max() <= 1e-6
    s.close()


def test_schur_system_matches_oracle(bab, orc):
    pb = syn.flatten(syn.make_scene("t", 5, 8, 300, 4.0, 8, arc_deg=50.0))
    s = bab.Solver(bab.default_options())
    s.set_problem(pb)
    S, rhs, cost = s.schur_system(1e4)
    So, rhso, costo = orc.schur_system(pb, 1e4, 0, pb.pt.shape[0])
    n = S.shape[0]
    # GPU: lower blocks valid, identity rows for fixed cameras; oracle: upper blocks, zero rows for fixed cameras
    L = np.tril(S)
    Lo = np.tril(So.T)
    for c in range(pb.cam.shape[0]):
        if pb.cam_fixed[c]:
            L[6 * c:6 * c + 6, 6 * c:6 * c + 6] = 0.0
    # block-lower vs element-lower: compare on the element-lower triangle
    assert np.abs(L - Lo).max() <= 1e-10 * np.abs(Lo).max()
    assert np.abs(rhs - rhso).max() <= 1e-10 * np.abs(rhso).max()
    assert abs(cost - costo) <= 1e-12 * costo
    s.close()


def test_schur_system_elimination_order(bab, orc):
    """Many camera groups with banded co-visibility: the device matrix lives in nested-dissection order (blocks of later-
    eliminated cameras are stored transposed); ba_schur_system must hand back the same S as the oracle."""
    pb = syn.flatten(syn.make_scene("t", 6, 150, 4000, 4.0, 16, arc_deg=120.0))
    s = bab.Solver(bab.default_options())
    s.set_problem(pb)
    st = s.reduced_system_stats()
    assert st["filled_tiles"] < st["lower_tiles"]
    S, rhs, cost = s.schur_system(1e4)
    So, rhso, costo = orc.schur_system(pb, 1e4, 0, pb.pt.shape[0])
    L = np.tril(S)
    Lo = np.tril(So.T)
    for c in range(pb.cam.shape[0]):
        if pb.cam_fixed[c]:
            L[6 * c:6 * c + 6, 6 * c:6 * c + 6] = 0.0
    assert np.abs(L - Lo).max() <= 1e-10 * np.abs(Lo).max()
    assert np.abs(rhs - rhso).max() <= 1e-10 * np.abs(rhso).max()
    s.close()


@pytest.mark.parametrize("tile", [64, 128])
@pytest.mark.parametrize("n", [6, 64, 130, 174, 500, 1194])
def test_dense_spd_solve(bab, n, tile):
    rng = np.random.default_rng(n)
    B = rng.normal(size=(n, n))
    A = B @ B.T + n * np.eye(n)
    b = rng.normal(size=n)
    x, ms = bab.capi.dense_spd_solve(A, b, tile_edge=tile)
    ref = np.linalg.solve(A, b)
    assert np.abs(x - ref).max() <= 1e-10 * np.abs(ref).max()


def _tile_pattern_matrix(n, pattern, seed):
    """SPD matrix whose 128x128 tile pattern follows pattern(i, j) (i >= j); diagonally dominant."""
    rng = np.random.default_rng(seed)
    T = (n + 127) // 128
    A = np.zeros((n, n))
    for i in range(T):
        for j in range(i + 1):
            if i == j or pattern(i, j, T):
                r0, r1, c0, c1 = i * 128, min(n, i * 128 + 128), j * 128, min(n, j * 128 + 128)
                A[r0:r1, c0:c1] = rng.normal(size=(r1 - r0, c1 - c0))
    A = np.tril(A)
    A = A + A.T
    A += np.diag(np.abs(A).sum(1) + 1.0)
    return A


@pytest.mark.parametrize("name,n,pattern", [
    ("band1", 1000, lambda i, j, T: i - j <= 1),
    ("band2", 1300, lambda i, j, T: i - j <= 2),
    ("arrow", 1100, lambda i, j, T: i == T - 1),                    # dense last block row: no fill
    ("arrow_first", 900, lambda i, j, T: j == 0),                   # dense first block column: fills everything
    ("blockdiag", 700, lambda i, j, T: False),
    ("scattered", 1500, lambda i, j, T: (7 * i + 3 * j) % 5 == 0),
])
@pytest.mark.parametrize("tile", [0, 64, 128])
def test_dense_spd_solve_tile_patterns(bab, name, n, pattern, tile):
    """The factorisation skips tiles that are structurally zero (tile-level symbolic fill); results must equal a
    dense solve for any pattern, including ones that fill in -- with 128-row tiles, with 64-row tiles (the choice for
    tile-sparse systems) and with the automatic choice."""
    A = _tile_pattern_matrix(n, pattern, seed=len(name) + n)
    b = np.random.default_rng(n).normal(size=n)
    x, ms = bab.capi.dense_spd_solve(A, b, tile_edge=tile)
    ref = np.linalg.solve(A, b)
    assert np.abs(x - ref).max() <= 1e-11 * np.abs(ref).max()


def test_tile_skipping_changes_nothing(bab):
    """Banded co-visibility (K4-shaped, scaled down): the LM trajectory with the tile-sparsity plan equals the one with
    every tile treated as filled up to summation order (the skipped terms are exact zeros; the plan only decides in which
    order a tile's contributions are added)."""
    pb = syn.flatten(syn.make_scene("band", 11, 260, 9000, 4.0, 16))
    opt = bab.default_options(max_num_iterations=6)
    runs = []
    for dense, tile in ((False, 128), (True, 128), (False, 64)):
        s = bab.Solver(opt)
        s.set_reduced_system_mode(dense_tiles=dense)
        s.set_reduced_system_tile(tile)
        s.set_problem(pb)
        sm = s.solve()
        st = s.reduced_system_stats()
        runs.append((np.array([it["cost"] for it in s.iterations()]), s.cameras(), s.points(), st))
        s.close()
    assert runs[0][3]["filled_tiles"] < runs[1][3]["filled_tiles"] == runs[1][3]["lower_tiles"]
    assert runs[0][0].shape == runs[1][0].shape
    assert np.abs(runs[0][0] - runs[1][0]).max() <= 1e-10 * runs[1][0].max()
    assert np.abs(runs[0][1] - runs[1][1]).max() <= 1e-8
    assert np.percentile(np.abs(runs[0][2] - runs[1][2]).max(axis=1), 99.0) <= 1e-8
    # 64-row tiles: another summation order of the same factorisation
    assert runs[2][0].shape == runs[1][0].shape and np.abs(runs[2][0] - runs[1][0]).max() <= 1e-10 * runs[1][0].max()
    assert np.abs(runs[2][1] - runs[1][1]).max() <= 1e-8


def test_dense_spd_not_positive_definite(bab):
    A = np.eye(8)
    A[3, 3] = -1.0
    with pytest.raises(bab.capi.BAError):
        bab.capi.dense_spd_solve(A, np.ones(8))


def test_solve_k1_trajectory(bab, orc, k1_small):
    pb = k1_small
    opt = bab.default_options(max_num_iterations=50)
    s = bab.Solver(opt)
    s.set_problem(pb)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=50))
    _traj_check(s.iterations(), oits)
    assert sm.termination_type == osm.termination_type
    assert sm.message == osm.message
    assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
    assert abs(sm.initial_cost - osm.initial_cost) <= 1e-12 * osm.initial_cost
    assert np.abs(s.cameras() - cam).max() <= 1e-6
    assert np.abs(s.points() - pt).max() <= 1e-6
    # re-run from the device-resident initial parameters: same answer
    s.reset_parameters()
    sm2 = s.solve()
    assert abs(sm2.final_cost - sm.final_cost) <= 1e-10 * sm.final_cost
    s.close()


def test_solve_k1_full(bab, orc):
    pb = syn.flatten(syn.config("K1"))
    opt = bab.default_options(max_num_iterations=100)
    s = bab.Solver(opt)
    s.set_problem(pb)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=100, num_threads=8))
    _traj_check(s.iterations(), oits)
    assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
    assert np.abs(s.cameras() - cam).max() <= 1e-6
    assert np.abs(s.points() - pt).max() <= 1e-6
    s.close()


def test_solve_k2_scaled(bab, orc):
    pb = syn.flatten(syn.config("K2", scale=0.3))
    opt = bab.default_options(max_num_iterations=30)
    s = bab.Solver(opt)
    s.set_problem(pb)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=30, num_threads=8))
    _traj_check(s.iterations(), oits)
    assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
    assert np.abs(s.cameras() - cam).max() <= 1e-6
    assert np.abs(s.points() - pt).max() <= 1e-6
    s.close()


def test_max_iterations_and_unobserved_blocks(bab, orc):
    """Landmarks / cameras without observations are not part of the problem and come back untouched."""
    import dataclasses
    pb = syn.flatten(syn.make_scene("t", 9, 7, 150, 3.0, 4, arc_deg=40.0))
    keep = (pb.obs_cam != 3) & (pb.obs_pt != 5) & (pb.obs_pt != 17)   # camera 3, landmarks 5 and 17 lose all observations
    pb = dataclasses.replace(pb, obs_cam=pb.obs_cam[keep].copy(), obs_pt=pb.obs_pt[keep].copy(), obs_uv=pb.obs_uv[keep].copy())
    opt = bab.default_options(max_num_iterations=3)
    s = bab.Solver(opt)
    s.set_problem(pb)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=3))
    _traj_check(s.iterations(), oits)
    assert sm.termination_type == bab.capi.NO_CONVERGENCE == osm.termination_type
    gcam, gpt = s.cameras(), s.points()
    assert np.array_equal(gcam[3], pb.cam[3]) and np.array_equal(gcam[0], pb.cam[0])
    unobs = np.bincount(pb.obs_pt, minlength=pb.pt.shape[0]) == 0
    assert unobs.any() and np.array_equal(gpt[unobs], pb.pt[unobs])
    assert np.abs(gcam - cam).max() <= 1e-6 and np.abs(gpt - pt).max() <= 1e-6
    s.close()


def test_pcg_tight_matches_dense(bab, orc, k1_small):
    """Schur-Jacobi PCG run to a tight tolerance reproduces the exact-solve (DENSE_SCHUR) trajectory."""
    pb = k1_small
    kw = dict(max_num_iterations=12, linear_solver=1, eta=1e-14, cg_r_tolerance=1e-13, max_linear_solver_iterations=2000)
    s = bab.Solver(bab.default_options(**kw))
    s.set_problem(pb)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=12))
    _traj_check(s.iterations(), oits)
    assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
    s.close()


def test_pcg_ceres_rule_matches_oracle_pcg(bab, orc, k1_small):
    """Ceres' ITERATIVE_SCHUR stopping rule (eta = 0.1): inexact steps, compared with the oracle's own PCG."""
    pb = k1_small
    kw = dict(max_num_iterations=15, linear_solver=1)
    s = bab.Solver(bab.default_options(**kw))
    s.set_problem(pb)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb, orc.default_options(**kw))
    its = s.iterations()
    _traj_check(its, oits, rtol=1e-6)
    assert [i["linear_solver_iterations"] for i in its] == [i["linear_solver_iterations"] for i in oits]
    s.close()


def test_pcg_matvec_modes(bab, orc):
    """ba_set_pcg_mode: the ordered matvec (camera-keyed fold, no floating-point atomics) is bit-reproducible run to run and
    is what the automatic mode picks for a problem of this size; the atomic variant follows the same trajectory."""
    pb = syn.flatten(syn.config("K2", scale=0.3))
    kw = dict(max_num_iterations=8, linear_solver=1)
    runs = {}
    for mode in (1, 1, 0, 2):
        s = bab.Solver(bab.default_options(**kw))
        s.set_pcg_mode(mode)
        s.set_problem(pb)
        s.solve()
        runs.setdefault(mode, []).append(([i["cost"] for i in s.iterations()], [i["linear_solver_iterations"] for i in s.iterations()],
                                          s.cameras(), s.points()))
        if mode == 1 and len(runs[1]) == 1:                 # and again on the SAME handle
            s.reset_parameters()
            s.solve()
            again = ([i["cost"] for i in s.iterations()], s.cameras(), s.points())
            assert again[0] == runs[1][0][0] and np.array_equal(again[1], runs[1][0][2]) and np.array_equal(again[2], runs[1][0][3])
        s.close()
    a, b = runs[1]
    assert a[0] == b[0] and np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])          # bit for bit across handles
    auto = runs[0][0]
    assert auto[0] == a[0] and np.array_equal(auto[2], a[2])                                   # automatic mode = ordered here
    at = runs[2][0]
    assert at[1] == a[1]
    assert np.abs(np.array(at[0]) - np.array(a[0])).max() <= 1e-9 * max(a[0])
    ocam, opt_, osm, oits = orc.solve(pb, orc.default_options(**kw))
    _traj_check([dict(iteration=i, step_is_successful=o["step_is_successful"], cost=c, trust_region_radius=o["trust_region_radius"])
                 for i, (c, o) in enumerate(zip(a[0], oits))], oits)


def test_pose_only_batch(bab, orc):
    batch = syn.make_pose_batch(seed=3, n_prob=64, n_pts=200)
    opt = bab.default_options(max_num_iterations=20)
    cam, sms, iters = bab.solve_pose_only_batch(batch, opt)
    ocam, osms, oiters = orc.solve_pose_only_batch(batch, orc.default_options(max_num_iterations=20, num_threads=8))
    for b in range(batch.cam.shape[0]):
        _traj_check(iters[b], oiters[b])
        assert sms[b].termination_type == osms[b].termination_type
        assert abs(sms[b].final_cost - osms[b].final_cost) <= 1e-8 * osms[b].final_cost
    assert np.abs(cam - ocam).max() <= 1e-6
    # masked observations are skipped (frame->isOutlier)
    mask = (np.random.default_rng(1).random(batch.uv.shape[0]) > 0.2).astype(np.uint8)
    cam2, _, it2 = bab.solve_pose_only_batch(batch, opt, mask=mask)
    ocam2, _, oit2 = orc.solve_pose_only_batch(batch, orc.default_options(max_num_iterations=20), mask=mask)
    assert np.abs(cam2 - ocam2).max() <= 1e-6
    # empty problems are legal
    import dataclasses
    empty = dataclasses.replace(batch, prob_ptr=np.zeros_like(batch.prob_ptr))
    cam3, sms3, _ = bab.solve_pose_only_batch(empty, opt)
    assert np.array_equal(cam3, batch.cam)


def test_prune_frames_bit_exact(bab, orc):
    rng = np.random.default_rng(5)
    sc = syn.config("K1", scale=0.2)
    frames = sc.pose_f32.shape[0]
    ptr = [0]
    xy, octv, xyz, mn, mx, has, out = [], [], [], [], [], [], []
    for f in range(frames):
        sel = np.nonzero(sc.obs_cam == f)[0]
        n = sel.shape[0] + 5
        ptr.append(ptr[-1] + n)
        X = np.concatenate([sc.pt_f32[sc.obs_pt[sel]], rng.normal(size=(5, 3)).astype(np.float32)])
        d = np.linalg.norm(X - sc.pose_f32[f, :3, 3], axis=1).astype(np.float32)
        xy.append(np.concatenate([sc.obs_uv[sel] + rng.normal(scale=2.0, size=(sel.shape[0], 2)).astype(np.float32),
                                  rng.uniform(0, 600, size=(5, 2)).astype(np.float32)]))
        octv.append(rng.integers(0, 6, size=n).astype(np.int32))
        xyz.append(X)
        mn.append((d * rng.uniform(0.3, 1.05, size=n)).astype(np.float32))
        mx.append((d * rng.uniform(0.95, 2.0, size=n)).astype(np.float32))
        has.append((rng.random(n) > 0.1).astype(np.uint8))
        out.append((rng.random(n) > 0.7).astype(np.uint8))
    cat = lambda l: np.concatenate(l)
    args = (cat(xy), cat(octv), cat(xyz), cat(mn), cat(mx), cat(has), cat(out))
    for consider in (True, False):
        got = bab.prune_frames(sc.pose_f32.reshape(-1, 16), sc.K_f32.reshape(-1, 9), np.array(ptr, np.int32), *args,
                               consider_outlier=consider)
        want = []
        for f in range(frames):
            a, b = ptr[f], ptr[f + 1]
            # the oracle inverts the pose itself; hand it the product's float32 inverse convention via the pose
            want.append(orc.prune_frame(sc.pose_f32[f], sc.K_f32[f], args[0][a:b], args[1][a:b], args[2][a:b], args[3][a:b],
                                        args[4][a:b], args[5][a:b], args[6][a:b], consider))
        want = np.concatenate(want)
        assert np.array_equal(got, want), int((got != want).sum())
    assert got.sum() > 0 and (got == 0).sum() > 0
