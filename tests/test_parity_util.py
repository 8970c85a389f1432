"""bundleadjustment_b200/parity.py (the gate bench.py applies before timing counts): host numpy logic, CPU test."""
import numpy as np

from bundleadjustment_b200 import parity, synthetic as syn


def _sol(orc, pb, **kw):
    cam, pt, sm, its = orc.solve(pb, orc.default_options(max_num_iterations=8, **kw))
    return {"costs": [i["cost"] for i in its], "final_cost": sm.final_cost, "cam": cam, "pt": pt, "radius": its[-1]["trust_region_radius"]}


def test_compare_accepts_equal_and_rejects_perturbed(orc):
    pb = syn.flatten(syn.make_scene("t", 5, 8, 300, 4.0, 8, arc_deg=50.0))
    a, b = _sol(orc, pb, num_threads=1), _sol(orc, pb, num_threads=4)
    r = parity.compare(pb, a, b)
    assert r["ok"] and r["max_cam_diff"] <= 1e-9 and r["pt_outside_cond_bound"] == 0
    bad = dict(a, pt=a["pt"] + 1e-4)
    r = parity.compare(pb, bad, b)
    assert not r["ok"] and r["pt_outside_cond_bound"] > 0
    bad = dict(a, costs=[c * (1 + 3e-6) for c in a["costs"]])
    assert not parity.compare(pb, bad, b)["ok"]
    bad = dict(a, costs=a["costs"][:-1])
    assert not parity.compare(pb, bad, b)["ok"]


def test_landmark_condition_matches_oracle_jacobians(orc):
    """the numpy pinhole Jacobian of parity.py against the oracle's (unrobustified: use_huber = 0)"""
    pb = syn.flatten(syn.make_scene("t", 6, 7, 120, 3.0, 4, arc_deg=40.0))
    ev = orc.evaluate(pb, orc.default_options(use_huber=0))
    Jp = ev["jac_pt"].reshape(-1, 2, 3)
    V = np.zeros((pb.pt.shape[0], 3, 3))
    np.add.at(V, pb.obs_pt, np.einsum('nki,nkj->nij', Jp, Jp))
    d = np.clip(np.einsum('nii->ni', V), 1e-6, 1e32) / 1e4
    V[:, 0, 0] += d[:, 0]; V[:, 1, 1] += d[:, 1]; V[:, 2, 2] += d[:, 2]
    w = np.linalg.eigvalsh(V)
    seen = np.nonzero(np.bincount(pb.obs_pt, minlength=pb.pt.shape[0]) > 0)[0]
    want = (w[:, 2] / w[:, 0])[seen]
    got = parity.landmark_condition(pb, pb.cam, pb.pt, seen, 1e4)
    assert np.abs(got - want).max() <= 1e-6 * want.max()
