"""CPU-only checks of the C-ABI library: it loads, exports every symbol include/ba_b200.h declares, its struct
layouts match the ctypes mirrors, its host-only helpers agree with the oracle, and compute entry points FAIL
LOUDLY without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ba_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ba_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(bab):
    syms = _declared_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(bab.capi.lib, s), f"{s} declared in include/ba_b200.h but not exported"
    assert sorted(bab.capi.EXPORTS) == syms
    assert bab.capi.lib.ba_version() == 100


def test_struct_layouts(bab, orc):
    # the header promises orc_options / ba_options share a layout (tests share initialisers)
    assert C.sizeof(bab.capi.Options) == C.sizeof(orc.Options) == 8 * 4 + 12 * 8
    assert C.sizeof(bab.capi.Iteration) == 4 * 4 + 8 * 8
    assert C.sizeof(bab.capi.Summary) == 4 * 4 + 5 * 8 + 128
    o = bab.default_options()
    assert o.max_num_iterations == 200 and o.num_threads == 4 and o.linear_solver == bab.capi.DENSE_SCHUR
    assert o.function_tolerance == 1e-6 and o.gradient_tolerance == 1e-10 and o.parameter_tolerance == 1e-8
    assert o.initial_trust_region_radius == 1e4 and o.min_relative_decrease == 1e-3
    assert o.huber_delta == np.sqrt(5.991) and o.use_huber == 1 and o.eta == 0.1
    oo = orc.default_options()
    for (f, _), (g, _) in zip(bab.capi.Options._fields_, orc.Options._fields_):
        # same slot, same default; ba_options.pad0 is orc_options.exploit_sparsity (0: the oracle factorises S densely, as Ceres does)
        assert f == g or (f, g) == ("pad0", "exploit_sparsity")
        assert getattr(o, f) == getattr(oo, g), f


def test_host_geometry_helpers_match_oracle(bab, orc):
    rng = np.random.default_rng(0)
    for _ in range(100):
        w = rng.normal(size=3) * rng.choice([1e-10, 1e-2, 1.0, 3.0])
        R = np.zeros(9)
        bab.capi.lib.ba_angle_axis_to_rotation(w.ctypes.data_as(C.c_void_p), R.ctypes.data_as(C.c_void_p))
        assert np.array_equal(R.reshape(3, 3).T, orc.angle_axis_to_rotation(w))
        w2 = np.zeros(3)
        bab.capi.lib.ba_rotation_to_angle_axis(R.ctypes.data_as(C.c_void_p), w2.ctypes.data_as(C.c_void_p))
        assert np.array_equal(w2, orc.rotation_to_angle_axis(R.reshape(3, 3).T))
        M = np.eye(4, dtype=np.float32)
        M[:3, :3] = orc.angle_axis_to_rotation(w).astype(np.float32)
        M[:3, 3] = rng.normal(size=3).astype(np.float32)
        assert np.array_equal(bab.capi.inverse4x4_f32(M), orc.inverse4x4_f32(M))
        cam, extr = bab.capi.pose_to_cam(M)
        ocam, oextr = orc.pose_to_cam(M)
        assert np.array_equal(cam, ocam) and np.array_equal(extr, oextr)
        assert np.array_equal(bab.capi.cam_to_pose(cam), orc.cam_to_pose(cam))


def test_partition_landmarks(bab):
    rng = np.random.default_rng(1)
    n_pt = 1000
    t = 2 + rng.geometric(0.4, size=n_pt)
    obs_pt = np.repeat(np.arange(n_pt, dtype=np.int32), t)
    for ranks in (1, 2, 3, 8):
        for quad in (False, True):
            cut = bab.capi.partition_landmarks(n_pt, obs_pt, ranks, quadratic=quad)
            assert cut[0] == 0 and cut[-1] == n_pt and np.all(np.diff(cut) >= 0)
            w = (t * (t + 1) // 2) if quad else t
            loads = np.array([w[a:b].sum() for a, b in zip(cut, cut[1:])])
            assert loads.max() <= loads.mean() * 1.05 + w.max()
    # ragged: more ranks than landmarks, empty input
    cut = bab.capi.partition_landmarks(2, np.array([0, 0, 1, 1], np.int32), 4)
    assert cut[0] == 0 and cut[-1] == 2 and np.all(np.diff(cut) >= 0)
    cut = bab.capi.partition_landmarks(0, np.zeros(0, np.int32), 2)
    assert cut.tolist() == [0, 0, 0]


def test_compute_entry_points_fail_without_gpu(bab):
    """No CPU fallback: on a box without a CUDA device ba_create reports BA_ERR_CUDA with a message."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(bab.capi.BAError) as e:
        bab.Solver(bab.default_options())
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)
    with pytest.raises(bab.capi.BAError):
        bab.capi.dense_spd_solve(np.eye(4), np.ones(4))
    from bundleadjustment_b200 import synthetic as syn
    with pytest.raises(bab.capi.BAError):
        bab.solve_pose_only_batch(syn.make_pose_batch(n_prob=2, n_pts=10))


def test_product_never_imports_the_oracle():
    """The product path may not route through oracle/ (or any CPU fallback): scan the package sources."""
    pkg = os.path.join(ROOT, "bundleadjustment_b200")
    for dirpath, _, files in os.walk(pkg):
        if "_build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "ba_oracle" not in text and "import oracle" not in text and "from oracle" not in text, os.path.join(dirpath, f)


def test_tile_elimination_order_band_and_dense():
    """Nested dissection of the Cholesky tile graph (host only): a permutation; shortens the dependency chain of banded
    co-visibility by a large factor; leaves a dense graph alone."""
    from bundleadjustment_b200 import capi
    for G, band in [(81, 2), (40, 1), (7, 1), (1, 0)]:
        adj = np.zeros((G, G), np.uint8)
        for i in range(G):
            adj[i, max(0, i - band):i + 1] = 1
        order, st = capi.tile_elimination_order(adj)
        assert sorted(order.tolist()) == list(range(G))
        assert st["depth_natural"] == 2 * G
        assert st["depth_ordered"] <= st["depth_natural"]
        if G >= 40:
            assert st["depth_ordered"] * 4 <= st["depth_natural"]
            assert st["filled_ordered"] <= 2 * st["filled_natural"]
    G = 30
    adj = np.tril(np.ones((G, G), np.uint8))
    order, st = capi.tile_elimination_order(adj)
    assert sorted(order.tolist()) == list(range(G))
    assert st["depth_ordered"] == st["depth_natural"] == 2 * G
    assert st["filled_ordered"] == st["filled_natural"] == G * (G + 1) // 2
    # two disconnected halves: independent chains
    adj = np.zeros((20, 20), np.uint8)
    for i in range(20):
        adj[i, i] = 1
        if i % 10:
            adj[i, i - 1] = 1
    order, st = capi.tile_elimination_order(adj)
    assert sorted(order.tolist()) == list(range(20))
    assert st["depth_ordered"] <= 20


def test_every_entry_point_is_documented_for_the_maintainer():
    """INTEGRATION.md section 2 maps EVERY entry point of include/ba_b200.h to the reference interface it replaces (or says that
    it has none); profiles/traffic.json only cites capture files that are committed."""
    import json
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "ba_b200.h")).read()
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    symbols = sorted(set(re.findall(r"\b(ba_[a-z0-9_]+)\s*\(", header)))
    assert len(symbols) >= 40
    assert [s for s in symbols if s not in doc] == []
    traffic = json.load(open(os.path.join(root, "profiles", "traffic.json")))
    for entry in traffic.values():
        cited = entry["source"].split(" ")[0]
        assert os.path.exists(os.path.join(root, cited)), cited
