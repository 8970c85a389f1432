"""End-to-end tests of the C++ drop-in optimizer classes (bundleadjustment_b200/host/Optimizer.{h,cpp}) on a mock
SceneMap: same entry points, same outer loops, same write-back and prune as the reference (Optimizer.cpp:216-277,
417-457, 500-610), compared with the oracle flow in tests/ref_flow.py.  State is truncated to float32 between outer
iterations on both sides (Optimizer.cpp:256-257, 265-267), so results agree to float32 rounding."""
import numpy as np
import pytest

import ref_flow as RF
from bundleadjustment_b200 import synthetic as syn

pytestmark = pytest.mark.gpu


def _close(a: RF.Graph, b: RF.Graph, frames=None):
    assert np.abs(a.pt - b.pt).max() <= 2e-6 * max(np.abs(b.pt).max(), 1.0)
    assert np.abs(a.pose - b.pose).max() <= 2e-6 * max(np.abs(b.pose).max(), 1.0)
    diff = int((a.kp_outlier != b.kp_outlier).sum())
    assert diff <= max(1, a.kp_outlier.shape[0] // 5000), diff      # chi-square knife edges after float32 rounding


def test_global_ba_three_outer_iterations(orc):
    """The pipeline's final global BA shape (BundleAdjustment.cpp:228-230: 3 outer x <= 100)."""
    g = RF.graph_from_scene(syn.config("K1", scale=0.2))
    want = RF.global_ba(orc, g, n_outer=3, max_it=100)
    got = RF.run_dropin(g, 0, 3, 100)
    assert got.info["status"] == 0
    _close(got, want)
    assert got.kp_outlier.sum() > 0                       # the 1 % gross outliers get pruned
    assert abs(got.info["final_cost"] - want.sm.final_cost) <= 1e-5 * want.sm.final_cost
    # frame 0 is the anchor: its pose only goes through the angle-axis round trip
    assert np.abs(got.pose[0] - g.pose[0]).max() < 1e-5


def test_global_ba_keyframe_call_and_erase_outliers(orc):
    """Per-keyframe call shape (BundleAdjustment.cpp:204-207: 1 outer x <= 10, eraseOutliers = true)."""
    g = RF.graph_from_scene(syn.config("K1", scale=0.15))
    g.kp_outlier[5] = 1                                   # an observation already flagged: skipped by the problem build
    want = RF.global_ba(orc, g, n_outer=1, max_it=10)
    got = RF.run_dropin(g, 0, 1, 10, erase=True, current=g.pose.shape[0] - 1)
    assert got.info["status"] == 0 and got.info["num_iterations"] <= 11
    _close(got, want)
    # eraseOutlier (SfMHelper.cpp:1079-1123): outlier keypoints lose their map point
    out = got.kp_outlier.astype(bool) & (g.kp_point >= 0)
    assert out.any() and np.all(got.kp_point[out] == -1)
    keep = ~got.kp_outlier.astype(bool)
    assert np.array_equal(got.kp_point[keep], g.kp_point[keep]) or (got.alive == 0).any()


def test_local_ba_window(orc):
    sc = syn.config("K1", scale=0.5)
    g = RF.graph_from_scene(sc)
    Nc = g.pose.shape[0]
    current = Nc - 1
    cov = [Nc - 2, Nc - 3, Nc - 4, Nc - 6]               # ordered covisibility list of the current frame
    want = RF.local_ba(orc, g, current, cov, n_outer=1, max_it=10)
    got = RF.run_dropin(g, 1, 1, 10, current=current, cov=cov)
    assert got.info["status"] == 0
    _close(got, want)
    local, fixed, pts = want.window
    assert fixed, "window has no fixed observers: test does not exercise point-only residuals"
    untouched = [f for f in range(Nc) if f not in local]
    assert np.array_equal(got.pose[untouched], g.pose[untouched])        # fixed / outside frames keep their pose
    outside = np.setdiff1d(np.arange(g.pt.shape[0]), pts)
    assert np.array_equal(got.pt[outside], g.pt[outside])


@pytest.mark.parametrize("mode", [2, 3])
def test_motion_only_pose_estimation(orc, mode):
    """SfMHelper::estimatePoseUsingBA (SfMHelper.cpp:106-117): 4 outer x <= 20 with prune(considerOutlier = true);
    mode 3 = the batched entry point, one launch for all frames."""
    sc = syn.config("K1", scale=0.12)
    g = RF.graph_from_scene(sc)
    rng = np.random.default_rng(1)
    g.kp_outlier[rng.random(g.kp_outlier.shape[0]) < 0.05] = 1     # outliers can become inliers again
    want = RF.motion_only(orc, g, 4, 20)
    got = RF.run_dropin(g, mode, 4, 20)
    assert got.info["status"] == 0
    _close(got, want)
    assert np.array_equal(got.pt, g.pt)                           # landmarks are constants here
