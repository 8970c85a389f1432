"""Loads tests/golden/*.npz (see tests/golden/make_golden.py) as synthetic.Problem + expected outputs."""
import glob
import os

import numpy as np

from bundleadjustment_b200.synthetic import Problem

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(n for n in (os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
               if not n.startswith("scipy_"))
# optimum pins produced by scipy.optimize.least_squares (tests/golden/make_scipy_pin.py): code neither restatement shares
SCIPY_CASES = sorted(os.path.splitext(os.path.basename(p))[0][len("scipy_"):] for p in glob.glob(os.path.join(GOLDEN_DIR, "scipy_*.npz")))


def load_scipy_pin(name):
    z = np.load(os.path.join(GOLDEN_DIR, "scipy_" + name + ".npz"))
    pb = Problem(cam=z["cam"], cam_fixed=z["cam_fixed"], fixed_extr=z["fixed_extr"], K=z["K"], pt=z["pt"],
                 obs_cam=z["obs_cam"], obs_pt=z["obs_pt"], obs_uv=z["obs_uv"])
    return pb, z


# options that let the LM loop run to the optimum instead of stopping at Ceres' default function tolerance
CONVERGE = dict(max_num_iterations=800, function_tolerance=1e-16, parameter_tolerance=1e-16, gradient_tolerance=1e-14)


def check_against_scipy_pin(z, cam, pt, final_cost):
    """cost 1e-8 relative everywhere (held to 1e-10: both sides sit at the same minimum); parameters 1e-6 where the
    optimum is unique (two fixed cameras; with one fixed camera monocular BA keeps a scale gauge freedom)."""
    c = float(z["opt_cost"])
    assert abs(final_cost - c) <= 1e-10 * c, (final_cost, c)
    if int(z["gauge_fixed"]):
        assert np.abs(cam - z["opt_cam"]).max() <= 1e-6, np.abs(cam - z["opt_cam"]).max()
        assert np.abs(pt - z["opt_pt"]).max() <= 1e-6, np.abs(pt - z["opt_pt"]).max()

FIELDS = ("iteration", "step_is_valid", "step_is_successful", "cost", "cost_change", "gradient_max_norm", "step_norm",
          "relative_decrease", "trust_region_radius")


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    pb = Problem(cam=z["cam"], cam_fixed=z["cam_fixed"], fixed_extr=z["fixed_extr"], K=z["K"], pt=z["pt"],
                 obs_cam=z["obs_cam"], obs_pt=z["obs_pt"], obs_uv=z["obs_uv"])
    n = z["it_cost"].shape[0]
    its = [{f: (int(z["it_" + f][i]) if f in FIELDS[:3] else float(z["it_" + f][i])) for f in FIELDS} for i in range(n)]
    return pb, z, its


def check_against_golden(z, gits, ev, its, cam, pt, final_cost, termination):
    """north_star tolerances: trajectory 1e-6 relative, final cost 1e-8 relative, parameters 1e-6."""
    for k in ("residuals", "jac_cam", "jac_pt"):
        assert np.abs(ev[k] - z[k]).max() <= 1e-11 * max(np.abs(z[k]).max(), 1.0), k
    assert abs(ev["cost"] - float(z["cost0"])) <= 1e-13 * float(z["cost0"])
    assert len(its) == len(gits), (len(its), len(gits))
    for a, b in zip(its, gits):
        assert a["iteration"] == b["iteration"] and a["step_is_successful"] == b["step_is_successful"]
        assert abs(a["cost"] - b["cost"]) <= 1e-6 * b["cost"], (a, b)
        assert abs(a["trust_region_radius"] - b["trust_region_radius"]) <= 1e-5 * b["trust_region_radius"]
        assert abs(a["step_norm"] - b["step_norm"]) <= 1e-5 * max(b["step_norm"], 1e-12)
    assert abs(final_cost - float(z["final_cost"])) <= 1e-8 * float(z["final_cost"])
    assert termination == int(z["termination"])
    assert np.abs(cam - z["out_cam"]).max() <= 1e-6 and np.abs(pt - z["out_pt"]).max() <= 1e-6
