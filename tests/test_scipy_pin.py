"""The optimum pinned by code neither the oracle nor the CUDA path shares: scipy.optimize.least_squares on the reference's
objective (tests/golden/make_scipy_pin.py -> tests/golden/scipy_*.npz).  Both the CPU oracle (run here on the CPU) and the
CUDA path (-m gpu) must converge to scipy's minimum: final cost 1e-8 relative, parameters 1e-6 on the gauge-fixed cases."""
import pytest

import golden_util as G

EXPECTED = {"tiny_huber", "tiny_huber_g2", "tiny_l2", "tiny_l2_g2", "small_huber", "small_huber_g2", "k1_quarter", "k1_quarter_g2"}


def test_scipy_pins_present():
    assert set(G.SCIPY_CASES) >= EXPECTED


@pytest.mark.parametrize("name", sorted(EXPECTED))
def test_oracle_converges_to_scipy_optimum(orc, name):
    pb, z = G.load_scipy_pin(name)
    cam, pt, sm, its = orc.solve(pb, orc.default_options(use_huber=int(z["use_huber"]), **G.CONVERGE))
    assert sm.termination_type == 0, sm.message           # CONVERGENCE, not the iteration cap
    G.check_against_scipy_pin(z, cam, pt, sm.final_cost)


@pytest.mark.gpu
@pytest.mark.parametrize("solver", [0, 1])
@pytest.mark.parametrize("name", sorted(EXPECTED))
def test_cuda_converges_to_scipy_optimum(bab, name, solver):
    """dense Schur and (tight) Schur-Jacobi PCG both land on scipy's minimum"""
    pb, z = G.load_scipy_pin(name)
    kw = dict(G.CONVERGE)
    if solver == 1:
        kw.update(linear_solver=1, eta=1e-14, cg_r_tolerance=1e-13, max_linear_solver_iterations=3000)
    s = bab.Solver(bab.default_options(use_huber=int(z["use_huber"]), **kw))
    s.set_problem(pb)
    sm = s.solve()
    assert sm.termination_type == 0, sm.message
    G.check_against_scipy_pin(z, s.cameras(), s.points(), sm.final_cost)
    s.close()
