"""Diagnostic (GPU box): how far does the Schur-Jacobi PCG, driven to a small relative residual, follow the dense-Cholesky LM
trajectory on K4-shaped scenes?  Prints per-iteration costs and CG iteration counts for several residual tolerances."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bundleadjustment_b200 as bab
from bundleadjustment_b200 import synthetic as syn

for scale in (0.1, 1.0):
    pb = syn.flatten(syn.config("K4", scale=scale))
    s = bab.Solver(bab.default_options(max_num_iterations=5)); s.set_problem(pb); s.solve()
    dense = [i["cost"] for i in s.iterations()]; rad = [i["trust_region_radius"] for i in s.iterations()]
    print(f"K4 x{scale} dense  costs", " ".join(f"{c:.9e}" for c in dense)); print("   radius", " ".join(f"{r:.3e}" for r in rad))
    if scale == 0.1:
        S, rhs, cost = s.schur_system(1e4)
        n = S.shape[0]
        A = np.tril(S) + np.tril(S, -1).T
        act = np.abs(np.diag(A)) > 0
        w = np.linalg.eigvalsh(A[np.ix_(act, act)])
        print(f"   S at radius 1e4: order {act.sum()}, eig min {w[0]:.3e} max {w[-1]:.3e} cond {w[-1] / w[0]:.3e}")
        S, rhs, cost = s.schur_system(1e7)
        A = np.tril(S) + np.tril(S, -1).T
        w = np.linalg.eigvalsh(A[np.ix_(act, act)])
        print(f"   S at radius 1e7: eig min {w[0]:.3e} max {w[-1]:.3e} cond {w[-1] / w[0]:.3e}")
    s.close()
    for tol in (1e-8, 1e-11, 1e-13):
        s = bab.Solver(bab.default_options(max_num_iterations=5, linear_solver=1, eta=-1.0, cg_r_tolerance=tol, max_linear_solver_iterations=20000))
        s.set_problem(pb); s.solve(); its = s.iterations()
        print(f"K4 x{scale} pcg r_tol {tol:g}: costs", " ".join(f"{i['cost']:.9e}" for i in its), "| CG", [i["linear_solver_iterations"] for i in its],
              "| max rel diff vs dense", max(abs(a["cost"] - b) / b for a, b in zip(its, dense)))
        s.close()
