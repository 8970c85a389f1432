"""Micro-benchmark of the reduced-system Cholesky (ba_dense_spd_solve) at the order of K4's S (n = 10200)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bundleadjustment_b200 as bab

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10200
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
rng = np.random.default_rng(0)
A = rng.standard_normal((n, n)) * 0.01
A = np.tril(A) + np.tril(A, -1).T
A[np.arange(n), np.arange(n)] += 1.0 + 0.01 * np.sqrt(n) * 4
b = rng.standard_normal(n)
x, ms = bab.capi.dense_spd_solve(A, b, repeats=reps)
r = np.abs(A @ x - b).max()
fl = n ** 3 / 3 + 2 * n * n
print(f"n={n} best {ms:.3f} ms  {fl / ms / 1e9:.2f} TFLOP/s  residual {r:.2e}")
