"""On-disk formats: BAL text, the flat binary dump and the reference's trajectory file
(BundleAdjustment.cpp:248-268).  C++ (bundleadjustment_b200/host/ProblemIO.cpp, driven through the test shim) and Python
(bundleadjustment_b200/problem_io.py) read each other's files bit for bit; the BAL camera model maps onto the reference's
pinhole; the GPU solves what it loads and checkpoints what it solved."""
import ctypes as C
import dataclasses

import numpy as np
import pytest

import ref_flow as RF
from bundleadjustment_b200 import problem_io as pio
from bundleadjustment_b200 import synthetic as syn

FIELDS = [f.name for f in dataclasses.fields(syn.Problem)]


def _shim():
    lib = C.CDLL(RF.build_shim())
    lib.ba_host_shim_convert.restype = C.c_int
    lib.ba_host_shim_trajectory.restype = C.c_int
    lib.ba_host_shim_solve_file.restype = C.c_int
    return lib


def _convert(lib, src, sk, dst, dk, fix_first=1):
    shape = (C.c_int32 * 4)()
    err = C.create_string_buffer(512)
    rc = lib.ba_host_shim_convert(str(src).encode(), sk, str(dst).encode(), dk, fix_first, shape, err, 512)
    return rc, list(shape), err.value.decode()


def _same(a: syn.Problem, b: syn.Problem):
    for k in FIELDS:
        x, y = getattr(a, k), getattr(b, k)
        assert x.dtype == y.dtype and x.shape == y.shape and np.array_equal(x, y), k


def _small():
    return syn.flatten(syn.make_scene("io", 7, 6, 120, 3.0, 5, arc_deg=40.0))


def test_binary_round_trip_both_ways(tmp_path):
    lib = _shim()
    pb = _small()
    dist = np.random.default_rng(0).normal(size=(pb.cam.shape[0], 2))
    for d in (None, dist):
        pio.write_binary(tmp_path / "a.bin", pb, d)
        rc, shape, err = _convert(lib, tmp_path / "a.bin", 1, tmp_path / "b.bin", 1)         # C++ reads and re-writes
        assert rc == 0, err
        assert shape[:3] == list(pb.shape)
        assert (tmp_path / "a.bin").read_bytes() == (tmp_path / "b.bin").read_bytes()
        got, gd = pio.read_binary(tmp_path / "b.bin")
        _same(got, pb)
        assert (gd is None) if d is None else np.array_equal(gd, d)


def test_bal_round_trip_and_model(tmp_path, orc):
    lib = _shim()
    pb = _small()
    pio.write_bal(tmp_path / "a.txt", pb)
    py, dist = pio.read_bal(tmp_path / "a.txt")
    rc, shape, err = _convert(lib, tmp_path / "a.txt", 0, tmp_path / "a.bin", 1)
    assert rc == 0, err
    cc, cd = pio.read_binary(tmp_path / "a.bin")
    _same(cc, py)                                                   # C++ reader == Python reader, every array, bit for bit
    assert np.array_equal(cd, dist) and shape == [*pb.shape, 0]
    assert np.array_equal(py.cam, pb.cam) and np.array_equal(py.pt, pb.pt)            # %.17g: doubles survive the text
    K = pb.K.reshape(-1, 9)
    assert np.array_equal(py.K[:, 0], -K[:, 0]) and np.array_equal(py.K[:, 8], K[:, 8]) and not py.K[:, [1, 2, 3, 5, 6, 7]].any()
    assert py.cam_fixed[0] == 1 and not py.cam_fixed[1:].any()
    # same model: BAL pixel = -(pixel - principal point); the cost of the loaded problem is the cost of the original
    free = dataclasses.replace(pb, cam_fixed=np.zeros_like(pb.cam_fixed))
    py_free, _ = pio.read_bal(tmp_path / "a.txt", fix_first_camera=False)
    a, b = orc.evaluate(free), orc.evaluate(py_free)
    assert abs(a["cost"] - b["cost"]) <= 1e-6 * a["cost"]
    assert np.abs(a["residuals"] + b["residuals"]).max() <= 1e-4      # float32 rounding of (uv - c)
    # and BAL's own formula, written out: P = R X + t, p = -P/P.z, pixel = f p
    R = syn._rodrigues(py_free.cam[:, :3])
    P = np.einsum("nij,nj->ni", R[py_free.obs_cam], py_free.pt[py_free.obs_pt]) + py_free.cam[py_free.obs_cam, 3:]
    f = -py_free.K[py_free.obs_cam, 0].astype(np.float64)
    pix = f[:, None] * (-P[:, :2] / P[:, 2:3])
    opt = orc.default_options(use_huber=0)
    raw = orc.evaluate(py_free, opt)["residuals"]
    assert np.abs(raw - (pix - py_free.obs_uv)).max() <= 1e-9
    # C++ writer -> Python reader
    rc, _, err = _convert(lib, tmp_path / "a.bin", 1, tmp_path / "b.txt", 0)
    assert rc == 0, err
    back, _ = pio.read_bal(tmp_path / "b.txt")
    _same(back, py)


def test_bal_known_answer_and_errors(tmp_path):
    lib = _shim()
    text = "2 1 2\n0 0 -1.5e+0 2.25\n1 0 3 -.5\n" + "\n".join(
        ["0.1", "-0.2", "0.3", "1", "2", "3", "500", "1e-7", "-2e-13", "0", "0", "0", "0", "0", "0", "4.5e2", "0", "0", "1", "2", "3"]) + "\n"
    (tmp_path / "k.txt").write_text(text)
    pb, dist = pio.read_bal(tmp_path / "k.txt", fix_first_camera=False)
    assert pb.shape == (2, 1, 2)
    assert pb.obs_uv.tolist() == [[-1.5, 2.25], [3.0, -0.5]] and pb.obs_cam.tolist() == [0, 1]
    assert pb.cam[0].tolist() == [0.1, -0.2, 0.3, 1, 2, 3] and pb.K[0, 0] == -500 and pb.K[1, 4] == -450
    assert dist.tolist() == [[1e-7, -2e-13], [0, 0]] and pb.pt.tolist() == [[1, 2, 3]]
    rc, shape, err = _convert(lib, tmp_path / "k.txt", 0, tmp_path / "k.bin", 1, fix_first=0)
    assert rc == 0 and shape == [2, 1, 2, 1]                            # one camera carries distortion
    _same(pio.read_binary(tmp_path / "k.bin")[0], pb)
    (tmp_path / "t.txt").write_text(text[:60])                           # truncated
    rc, _, err = _convert(lib, tmp_path / "t.txt", 0, tmp_path / "t.bin", 1)
    assert rc != 0 and "truncated" in err
    with pytest.raises(ValueError, match="truncated"):
        pio.read_bal(tmp_path / "t.txt")
    (tmp_path / "r.txt").write_text(text.replace("1 0 3 -.5", "2 0 3 -.5"))
    rc, _, err = _convert(lib, tmp_path / "r.txt", 0, tmp_path / "r.bin", 1)
    assert rc != 0 and "out of range" in err
    with pytest.raises(ValueError, match="out of range"):
        pio.read_bal(tmp_path / "r.txt")
    (tmp_path / "g.bin").write_bytes(b"not a dump at all")
    rc, _, err = _convert(lib, tmp_path / "g.bin", 1, tmp_path / "g2.bin", 1)
    assert rc != 0 and "BAB200P" in err
    with pytest.raises(ValueError, match="BAB200P"):
        pio.read_binary(tmp_path / "g.bin")
    skew = dataclasses.replace(_small())
    skew.K = skew.K.copy(); skew.K.reshape(-1, 9)[2, 1] = 0.5
    pio.write_binary(tmp_path / "s.bin", skew)
    rc, _, err = _convert(lib, tmp_path / "s.bin", 1, tmp_path / "s.txt", 0)
    assert rc != 0 and "cannot express" in err
    with pytest.raises(ValueError, match="cannot express"):
        pio.write_bal(tmp_path / "s.txt", skew)


def test_trajectory_file(tmp_path):
    """BundleAdjustment.cpp:251-264: keyframes in ID order, `timestamp tx ty tz qx qy qz qw`, fixed, 4 decimals."""
    lib = _shim()
    rng = np.random.default_rng(3)
    n = 12
    poses = np.tile(np.eye(4, dtype=np.float32), (n, 1, 1))
    w = rng.normal(scale=1.5, size=(n, 3))
    w[0] = 0                                                   # identity
    w[1] = [0, 0, np.pi]                                       # trace < 0: the largest-diagonal branches
    w[2] = [np.pi, 0, 0]
    w[3] = [0, np.pi - 1e-3, 0]
    poses[:, :3, :3] = syn._rodrigues(w).astype(np.float32)
    poses[:, :3, 3] = rng.normal(scale=3.0, size=(n, 3)).astype(np.float32)
    poses[0, :3, 3] = 0
    ts = 1305031102.175304 + 0.0333 * np.arange(n)
    ids = np.arange(n, dtype=np.int32)
    pio.write_trajectory_tum(tmp_path / "py", ts, poses)
    perm = rng.permutation(n)                                  # inserted out of order: SceneMap orders keyframes by ID
    err = C.create_string_buffer(256)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    p2, t2, i2 = np.ascontiguousarray(poses[perm]), np.ascontiguousarray(ts[perm]), np.ascontiguousarray(ids[perm])
    assert lib.ba_host_shim_trajectory(str(tmp_path / "cc").encode(), n, P(i2), P(t2), P(p2), err, 256) == 0, err.value
    a = (tmp_path / "py_estimatedPoses.txt").read_text()
    b = (tmp_path / "cc_estimatedPoses.txt").read_text()
    assert a == b
    lines = a.splitlines()
    assert lines[0] == "1305031102.1753 0.0000 0.0000 0.0000 0.0000 0.0000 0.0000 1.0000"
    q1 = [float(x) for x in lines[1].split()[4:]]
    assert abs(abs(q1[2]) - 1) < 1e-4 and abs(q1[3]) < 1e-4
    for ln, Pm in zip(lines, poses):
        v = [float(x) for x in ln.split()]
        x, y, z, ww = v[4:]
        assert abs(x * x + y * y + z * z + ww * ww - 1) < 1e-3
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * ww), 2 * (x * z + y * ww)],
                      [2 * (x * y + z * ww), 1 - 2 * (x * x + z * z), 2 * (y * z - x * ww)],
                      [2 * (x * z - y * ww), 2 * (y * z + x * ww), 1 - 2 * (x * x + y * y)]])
        assert np.abs(R - Pm[:3, :3]).max() < 2e-3             # 4 decimals


@pytest.mark.gpu
def test_load_solve_checkpoint(tmp_path, bab, orc):
    """BAL text -> C++ reader -> binary -> C++ load/solve/checkpoint on the GPU; the oracle solves the same file."""
    lib = _shim()
    pb = syn.flatten(syn.config("K1", scale=0.2))
    pio.write_bal(tmp_path / "p.txt", pb)
    rc, shape, err = _convert(lib, tmp_path / "p.txt", 0, tmp_path / "p.bin", 1)
    assert rc == 0, err
    info = np.zeros(4)
    ebuf = C.create_string_buffer(512)
    rc = lib.ba_host_shim_solve_file(str(tmp_path / "p.bin").encode(), str(tmp_path / "out.bin").encode(), 15, 0,
                                     info.ctypes.data_as(C.c_void_p), ebuf, 512)
    assert rc == 0, ebuf.value
    loaded, _ = pio.read_binary(tmp_path / "p.bin")
    cam, pt, osm, oits = orc.solve(loaded, orc.default_options(max_num_iterations=15))
    assert abs(info[0] - osm.initial_cost) <= 1e-9 * osm.initial_cost
    assert abs(info[1] - osm.final_cost) <= 1e-8 * osm.final_cost
    out, _ = pio.read_binary(tmp_path / "out.bin")
    assert np.abs(out.cam - cam).max() <= 1e-6 and np.abs(out.pt - pt).max() <= 1e-6
    assert np.array_equal(out.cam[0], loaded.cam[0])                # the anchor's entry is kept
    for k in ("cam_fixed", "fixed_extr", "K", "obs_cam", "obs_pt", "obs_uv"):
        assert np.array_equal(getattr(out, k), getattr(loaded, k))
    # the checkpoint is a problem: resuming from it starts at the cost the first solve ended on, through the Python binding
    s = bab.Solver(bab.default_options(max_num_iterations=5))
    s.set_problem(out)
    sm = s.solve()
    assert abs(sm.initial_cost - info[1]) <= 1e-9 * info[1]
    s.close()
