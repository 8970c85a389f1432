"""Edge cases of the CUDA path through the C ABI: bad inputs fail loudly, degenerate problems behave like the oracle,
long tracks / duplicate observations / multi-unit Schur blocks agree with the oracle, and the full-size K4 workload
satisfies size-independent properties (the oracle would need minutes there)."""
import dataclasses

import numpy as np
import pytest

from bundleadjustment_b200 import synthetic as syn

pytestmark = pytest.mark.gpu


def _traj(its):
    return np.array([i["cost"] for i in its])


def test_bad_inputs_fail_loudly(bab):
    pb = syn.flatten(syn.make_scene("t", 3, 5, 60, 3.0, 4, arc_deg=40.0))
    s = bab.Solver(bab.default_options())
    with pytest.raises(bab.capi.BAError):
        s.solve()                                   # no problem set
    bad = dataclasses.replace(pb, obs_cam=pb.obs_cam.copy())
    bad.obs_cam[7] = pb.cam.shape[0]                # camera index out of range
    with pytest.raises(bab.capi.BAError, match="out of range"):
        s.set_problem(bad)
    bad = dataclasses.replace(pb, obs_pt=pb.obs_pt.copy())
    bad.obs_pt[0] = -1
    with pytest.raises(bab.capi.BAError, match="out of range"):
        s.set_problem(bad)
    s.set_problem(pb)                               # the handle is still usable afterwards
    assert s.solve().final_cost > 0
    s.close()


def test_no_observations(bab):
    """An empty observation list is a legal (trivial) problem: zero cost, parameters untouched, no crash."""
    pb = syn.flatten(syn.make_scene("t", 3, 4, 30, 3.0, 4, arc_deg=40.0))
    empty = dataclasses.replace(pb, obs_cam=pb.obs_cam[:0].copy(), obs_pt=pb.obs_pt[:0].copy(), obs_uv=pb.obs_uv[:0].copy())
    for solver in (0, 1):
        s = bab.Solver(bab.default_options(max_num_iterations=5, linear_solver=solver))
        s.set_problem(empty)
        sm = s.solve()
        assert sm.initial_cost == 0.0 and sm.final_cost == 0.0
        assert np.array_equal(s.cameras(), pb.cam) and np.array_equal(s.points(), pb.pt)
        s.close()


def test_long_tracks_and_multi_unit_blocks(bab, orc):
    """Landmarks seen by up to 128 cameras (tracks longer than half a tile window) and camera pairs sharing more than 256
    landmarks (Schur blocks reduced over several units): dense and PCG paths against the oracle."""
    sc = syn.make_scene("long", 21, 140, 1500, 60.0, 128, arc_deg=200.0, ring_radius=8.0, cloud_radius=2.0)
    pb = syn.flatten(sc)
    track = np.bincount(pb.obs_pt, minlength=pb.pt.shape[0])
    assert track.max() >= 100
    opt = bab.default_options(max_num_iterations=6)
    s = bab.Solver(opt)
    s.set_problem(pb)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb, orc.default_options(max_num_iterations=6))
    c, co = _traj(s.iterations()), _traj(oits)
    assert c.shape == co.shape and np.abs(c - co).max() <= 1e-6 * co.max()
    assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
    assert np.abs(s.cameras() - cam).max() <= 1e-6
    s.close()
    kw = dict(max_num_iterations=6, linear_solver=1, eta=1e-14, cg_r_tolerance=1e-13, max_linear_solver_iterations=3000)
    s = bab.Solver(bab.default_options(**kw))
    s.set_problem(pb)
    sm2 = s.solve()
    assert abs(sm2.final_cost - osm.final_cost) <= 1e-7 * osm.final_cost
    s.close()


def test_duplicate_observations(bab, orc):
    """The same (camera, landmark) pair observed twice (two keypoints matched to one map point) is two residual blocks."""
    pb = syn.flatten(syn.make_scene("dup", 8, 9, 200, 4.0, 8, arc_deg=50.0))
    rng = np.random.default_rng(1)
    idx = rng.choice(pb.obs_cam.shape[0], 40, replace=False)
    uv2 = pb.obs_uv[idx] + rng.normal(scale=0.3, size=(40, 2)).astype(np.float32)
    pb2 = dataclasses.replace(pb, obs_cam=np.concatenate([pb.obs_cam, pb.obs_cam[idx]]),
                              obs_pt=np.concatenate([pb.obs_pt, pb.obs_pt[idx]]),
                              obs_uv=np.ascontiguousarray(np.concatenate([pb.obs_uv, uv2])))
    s = bab.Solver(bab.default_options(max_num_iterations=8))
    s.set_problem(pb2)
    sm = s.solve()
    cam, pt, osm, oits = orc.solve(pb2, orc.default_options(max_num_iterations=8))
    c, co = _traj(s.iterations()), _traj(oits)
    assert c.shape == co.shape and np.abs(c - co).max() <= 1e-6 * co.max()
    assert abs(sm.final_cost - osm.final_cost) <= 1e-8 * osm.final_cost
    assert np.abs(s.cameras() - cam).max() <= 1e-6 and np.abs(s.points() - pt).max() <= 1e-6
    s.close()


def test_k4_full_size_properties(bab):
    """BASELINE's K4 at full size (1 700 cameras / 500 k landmarks / 2 M observations), properties that need no oracle:
    the reported initial cost is the cost ba_evaluate computes, accepted steps decrease the cost monotonically, a second
    solve from the stored initial parameters reproduces the trajectory bit for bit (every reduction is ordered), the
    PCG path driven to a 1e-12 residual follows the dense path (to the accuracy the conditioning of S allows: ten LM
    iterations of this unconverged problem agree to 1e-4 relative), and re-solving from the solution changes nothing much."""
    pb = syn.flatten(syn.config("K4"))
    s = bab.Solver(bab.default_options(max_num_iterations=10))
    s.set_problem(pb)
    ev = s.evaluate(jacobians=False, gradient=False)
    sm = s.solve()
    its = s.iterations()
    assert abs(sm.initial_cost - ev["cost"]) <= 1e-12 * ev["cost"]
    acc = [i["cost"] for i in its if i["step_is_successful"] or i["iteration"] == 0]
    assert all(b <= a for a, b in zip(acc, acc[1:])) and sm.final_cost < 0.05 * sm.initial_cost
    st = s.reduced_system_stats()
    assert st["filled_tiles"] * 4 < st["lower_tiles"]          # sliding-window co-visibility: most tiles stay empty
    cam1, pt1 = s.cameras(), s.points()
    s.reset_parameters()
    sm_b = s.solve()
    assert np.array_equal(_traj(s.iterations()), _traj(its))
    assert np.array_equal(s.cameras(), cam1) and np.array_equal(s.points(), pt1)
    s.close()
    kw = dict(max_num_iterations=10, linear_solver=1, eta=1e-14, cg_r_tolerance=1e-12, max_linear_solver_iterations=5000)
    s = bab.Solver(bab.default_options(**kw))
    s.set_problem(pb)
    sm_p = s.solve()
    assert abs(sm_p.final_cost - sm.final_cost) <= 5e-4 * sm.final_cost
    assert [i["step_is_successful"] for i in s.iterations()] == [i["step_is_successful"] for i in its]
    s.close()
    pb2 = dataclasses.replace(pb, cam=cam1, pt=pt1)
    s = bab.Solver(bab.default_options(max_num_iterations=3))
    s.set_problem(pb2)
    sm2 = s.solve()
    assert abs(sm2.initial_cost - sm_b.final_cost) <= 1e-9 * sm_b.final_cost
    assert sm2.final_cost <= sm2.initial_cost and sm2.final_cost >= 0.98 * sm2.initial_cost
    s.close()
