"""CPU-only checks of the C++ drop-in classes: the shim builds against the mock data model, and without a GPU the
entry points fail LOUDLY (status != 0, map untouched) instead of falling back to a CPU path."""
import numpy as np
import pytest

import ref_flow as RF
from bundleadjustment_b200 import synthetic as syn


def test_dropin_builds_and_fails_loudly_without_gpu(capfd):
    import torch
    g = RF.graph_from_scene(syn.config("K1", scale=0.1))
    got = RF.run_dropin(g, 0, 1, 5)
    if torch.cuda.is_available():
        assert got.info["status"] == 0
        return
    assert got.info["status"] != 0
    assert np.array_equal(got.pose, g.pose) and np.array_equal(got.pt, g.pt) and np.array_equal(got.kp_outlier, g.kp_outlier)
    assert "ba_b200" in capfd.readouterr().err
    for mode in (2, 3):
        got = RF.run_dropin(g, mode, 1, 5)
        assert got.info["status"] != 0 and np.array_equal(got.pose, g.pose)


def test_oracle_flow_outer_iterations(orc):
    """The oracle-side restatement of the entry point itself: prune between outer iterations changes the residual set,
    the state is float32 after every outer iteration, the anchor frame only sees an angle-axis round trip."""
    g = RF.graph_from_scene(syn.config("K1", scale=0.1))
    a = RF.global_ba(orc, g, n_outer=1, max_it=50)
    b = RF.global_ba(orc, g, n_outer=3, max_it=50)
    assert a.kp_outlier.sum() > 0 and b.kp_outlier.sum() >= a.kp_outlier.sum() - 2
    assert b.sm.initial_cost < a.sm.initial_cost          # outliers pruned after iteration 1 leave the problem
    assert np.abs(a.pose[0] - g.pose[0]).max() < 1e-5
    assert a.pt.dtype == np.float32 and a.pose.dtype == np.float32
    local, fixed, pts = RF.local_window(g, g.pose.shape[0] - 1, [0, 1])
    assert local[-1] == g.pose.shape[0] - 1 and len(set(local) & set(fixed)) == 0
