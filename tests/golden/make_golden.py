"""Generates tests/golden/*.npz -- small problems with their expected LM trajectories.

PARITY UNPINNED caveat: the reference's BA path cannot be executed here (its arithmetic is Ceres/Eigen, absent from
/root/reference and from this image; the reference ships no golden vectors), so these fixtures are produced by the
repo's own two independent CPU restatements and only record a case where BOTH agree: the C++ oracle
(oracle/ba_oracle.cpp) and the numpy mini-oracle (tests/mini_oracle.py: complex-step dense Jacobian + dense normal
equations).  Run from the repo root:   python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import mini_oracle                                   # noqa: E402
from bundleadjustment_b200 import synthetic as syn   # noqa: E402
from oracle import ba_oracle as orc                  # noqa: E402

CASES = {
    # name: (seed, cams, landmarks, mean track, window, huber, max_iter)
    "tiny_huber": (101, 4, 24, 3.0, 4, 1, 30),
    "tiny_l2": (102, 5, 30, 3.0, 4, 0, 30),
    "small_huber": (103, 6, 48, 3.5, 4, 1, 8),
}
FIELDS = ("iteration", "step_is_valid", "step_is_successful", "cost", "cost_change", "gradient_max_norm", "step_norm",
          "relative_decrease", "trust_region_radius")


def main():
    for name, (seed, nc, npt, track, win, huber, max_it) in CASES.items():
        pb = syn.flatten(syn.make_scene(name, seed, nc, npt, track, win, arc_deg=40.0))
        cam, pt, sm, its = orc.solve(pb, orc.default_options(max_num_iterations=max_it, use_huber=huber))
        mcam, mpt, term, mits = mini_oracle.solve(pb, max_it, use_huber=bool(huber))
        assert len(its) == len(mits), (name, len(its), len(mits))
        for a, b in zip(its, mits):
            assert a["step_is_successful"] == b["step_is_successful"]
            assert abs(a["cost"] - b["cost"]) <= 1e-8 * b["cost"], (name, a, b)
        assert np.abs(cam - mcam).max() < 1e-7 and np.abs(pt - mpt).max() < 1e-7
        ev = orc.evaluate(pb, orc.default_options(use_huber=huber))
        np.savez_compressed(
            os.path.join(HERE, f"{name}.npz"),
            cam=pb.cam, cam_fixed=pb.cam_fixed, fixed_extr=pb.fixed_extr, K=pb.K, pt=pb.pt, obs_cam=pb.obs_cam,
            obs_pt=pb.obs_pt, obs_uv=pb.obs_uv, use_huber=np.int32(huber), max_num_iterations=np.int32(max_it),
            out_cam=cam, out_pt=pt, residuals=ev["residuals"], jac_cam=ev["jac_cam"], jac_pt=ev["jac_pt"],
            cost0=np.float64(ev["cost"]), termination=np.int32(sm.termination_type), final_cost=np.float64(sm.final_cost),
            **{f"it_{f}": np.array([i[f] for i in its], dtype=np.float64) for f in FIELDS})
        print(name, "iterations", len(its), "cost", its[0]["cost"], "->", sm.final_cost, sm.message.decode())
    json.dump({"generator": "tests/golden/make_golden.py", "sources": ["oracle/ba_oracle.cpp", "tests/mini_oracle.py"],
               "cases": {k: list(v) for k, v in CASES.items()}}, open(os.path.join(HERE, "MANIFEST.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
