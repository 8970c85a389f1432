#!/usr/bin/env python3
"""A/B of the experimental ring variant of the PCG matvec (BA_B200_MATVEC_RING=1, csrc/ba_pcg.cu) against the measured kernel,
on a GPU box:

    timeout 300 python tools/check_matvec_ring.py [workload] [scale]          # default K4 x0.2

Each variant runs in its own process (the switch is read once per process): PCG solve of the same problem, cost trajectory
and device time printed; exits non-zero if the trajectories differ by more than 1e-9 relative (the f64 reductions of the
matvec are unordered, so bitwise equality is not expected)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import json, sys
sys.path.insert(0, %r)
import bundleadjustment_b200 as bab
from bundleadjustment_b200 import synthetic as syn
pb = syn.flatten(syn.config(sys.argv[1], scale=float(sys.argv[2])))
opt = bab.default_options(max_num_iterations=8, linear_solver=bab.capi.ITERATIVE_SCHUR)
s = bab.Solver(opt, device=0)
s.set_problem(pb)
s.solve()
s.reset_parameters()
sm = s.solve()
print(json.dumps({"costs": [i["cost"] for i in s.iterations()], "ms": sm.seconds_total * 1e3}))
s.close()
""" % ROOT


def run(ring: bool, workload: str, scale: str) -> dict:
    env = dict(os.environ, BA_B200_MATVEC_RING="1" if ring else "0")
    out = subprocess.run([sys.executable, "-c", CHILD, workload, scale], env=env, capture_output=True, text=True, timeout=280)
    if out.returncode != 0:
        raise SystemExit(f"variant ring={ring} failed:\n{out.stderr[-2000:]}")
    return json.loads(out.stdout.strip().splitlines()[-1])


def main() -> int:
    workload = sys.argv[1] if len(sys.argv) > 1 else "K4"
    scale = sys.argv[2] if len(sys.argv) > 2 else "0.2"
    a, b = run(False, workload, scale), run(True, workload, scale)
    worst = max(abs(x - y) / abs(x) for x, y in zip(a["costs"], b["costs"])) if len(a["costs"]) == len(b["costs"]) else float("inf")
    print(json.dumps({"workload": workload, "scale": scale, "iterations": len(a["costs"]), "max_cost_rel_diff": worst,
                      "ms_measured_kernel": a["ms"], "ms_ring_kernel": b["ms"]}))
    return 0 if worst <= 1e-9 else 1


if __name__ == "__main__":
    sys.exit(main())
