"""ctypes binding of oracle/_build/libba_oracle.so -- TEST INFRASTRUCTURE ONLY (see ba_oracle.h).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
PARITY UNPINNED: restates Ceres (absent) from its published algorithm; see the header of ba_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libba_oracle.so")

DENSE_SCHUR, ITERATIVE_SCHUR = 0, 1
CONVERGENCE, NO_CONVERGENCE, FAILURE = 0, 1, 2


class Options(C.Structure):
    _fields_ = [
        ("max_num_iterations", C.c_int32), ("num_threads", C.c_int32), ("linear_solver", C.c_int32),
        ("jacobi_scaling", C.c_int32), ("max_num_consecutive_invalid_steps", C.c_int32),
        ("max_linear_solver_iterations", C.c_int32), ("use_huber", C.c_int32), ("exploit_sparsity", C.c_int32),
        ("function_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
        ("parameter_tolerance", C.c_double), ("initial_trust_region_radius", C.c_double),
        ("max_trust_region_radius", C.c_double), ("min_trust_region_radius", C.c_double),
        ("min_relative_decrease", C.c_double), ("min_lm_diagonal", C.c_double),
        ("max_lm_diagonal", C.c_double), ("huber_delta", C.c_double), ("eta", C.c_double),
        ("cg_r_tolerance", C.c_double),
    ]


class Iteration(C.Structure):
    _fields_ = [
        ("iteration", C.c_int32), ("step_is_valid", C.c_int32), ("step_is_successful", C.c_int32),
        ("linear_solver_iterations", C.c_int32), ("cost", C.c_double), ("cost_change", C.c_double),
        ("gradient_max_norm", C.c_double), ("gradient_norm", C.c_double), ("step_norm", C.c_double),
        ("relative_decrease", C.c_double), ("trust_region_radius", C.c_double), ("eta", C.c_double),
    ]


class Summary(C.Structure):
    _fields_ = [
        ("termination_type", C.c_int32), ("num_iterations", C.c_int32),
        ("num_successful_steps", C.c_int32), ("num_unsuccessful_steps", C.c_int32),
        ("initial_cost", C.c_double), ("final_cost", C.c_double), ("seconds_total", C.c_double),
        ("seconds_linear_solver", C.c_double), ("seconds_evaluate", C.c_double), ("message", C.c_char * 128),
    ]


def build(force: bool = False) -> str:
    """Compile the oracle (g++ -O3 -fopenmp).  Building the checker is not using it."""
    src = os.path.join(_HERE, "ba_oracle.cpp")
    hdr = os.path.join(_HERE, "ba_oracle.h")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["make", "-C", _HERE, "-s"], env={**os.environ, "CXX": "/usr/bin/g++"})
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
    return _lib


def _p(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def default_options(**kw) -> Options:
    o = Options()
    lib().orc_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


def angle_axis_to_rotation(w):
    w = np.ascontiguousarray(w, np.float64)
    R = np.zeros(9)
    lib().orc_angle_axis_to_rotation(_p(w, C.c_double), _p(R, C.c_double))
    return R.reshape(3, 3).T.copy()        # column-major -> numpy [row, col]


def rotation_to_angle_axis(R):
    Rc = np.ascontiguousarray(np.asarray(R, np.float64).T).reshape(9)
    w = np.zeros(3)
    lib().orc_rotation_to_angle_axis(_p(Rc, C.c_double), _p(w, C.c_double))
    return w


def inverse4x4_f32(M):
    M = np.ascontiguousarray(M, np.float32).reshape(16)
    out = np.zeros(16, np.float32)
    lib().orc_inverse4x4_f32(_p(M, C.c_float), _p(out, C.c_float))
    return out.reshape(4, 4)


def eval_cam_point(cam, pt, uv, K, jets=False):
    cam = np.ascontiguousarray(cam, np.float64); pt = np.ascontiguousarray(pt, np.float64)
    uv = np.ascontiguousarray(uv, np.float32); K = np.ascontiguousarray(K, np.float32).reshape(9)
    r, Jc, Jp = np.zeros(2), np.zeros(12), np.zeros(6)
    lib().orc_eval_cam_point(_p(cam, C.c_double), _p(pt, C.c_double), _p(uv, C.c_float), _p(K, C.c_float),
                             _p(r, C.c_double), _p(Jc, C.c_double), _p(Jp, C.c_double), int(jets))
    return r, Jc.reshape(2, 6), Jp.reshape(2, 3)


def eval_point_only(pt, uv, E, K, jets=False):
    pt = np.ascontiguousarray(pt, np.float64); uv = np.ascontiguousarray(uv, np.float32)
    E = np.ascontiguousarray(E, np.float32).reshape(16); K = np.ascontiguousarray(K, np.float32).reshape(9)
    r, Jp = np.zeros(2), np.zeros(6)
    lib().orc_eval_point_only(_p(pt, C.c_double), _p(uv, C.c_float), _p(E, C.c_float), _p(K, C.c_float),
                              _p(r, C.c_double), _p(Jp, C.c_double), int(jets))
    return r, Jp.reshape(2, 3)


def eval_pose_only(cam, Xh, uv, K, jets=False):
    cam = np.ascontiguousarray(cam, np.float64); Xh = np.ascontiguousarray(Xh, np.float32)
    uv = np.ascontiguousarray(uv, np.float32); K = np.ascontiguousarray(K, np.float32).reshape(9)
    r, Jc = np.zeros(2), np.zeros(12)
    lib().orc_eval_pose_only(_p(cam, C.c_double), _p(Xh, C.c_float), _p(uv, C.c_float), _p(K, C.c_float),
                             _p(r, C.c_double), _p(Jc, C.c_double), int(jets))
    return r, Jc.reshape(2, 6)


def huber(s, a):
    rho = np.zeros(3)
    f = lib().orc_huber
    f.argtypes = [C.c_double, C.c_double, C.POINTER(C.c_double)]
    f(float(s), float(a), _p(rho, C.c_double))
    return rho


def _prob_args(pb):
    Nc, Np, No = pb.cam.shape[0], pb.pt.shape[0], pb.obs_cam.shape[0]
    return Nc, Np, No


def evaluate(pb, opt=None, cam=None, pt=None):
    """Robustified residuals / Jacobians / gradient at (cam, pt) (defaults: the problem's own)."""
    opt = opt or default_options()
    Nc, Np, No = _prob_args(pb)
    cam = np.ascontiguousarray(pb.cam if cam is None else cam, np.float64)
    pt = np.ascontiguousarray(pb.pt if pt is None else pt, np.float64)
    cost = C.c_double(0)
    res, Jc, Jp = np.zeros((No, 2)), np.zeros((No, 2, 6)), np.zeros((No, 2, 3))
    gc, gp = np.zeros((Nc, 6)), np.zeros((Np, 3))
    rc = lib().orc_evaluate(C.byref(opt), Nc, _p(cam, C.c_double), _p(pb.cam_fixed, C.c_uint8),
                            _p(pb.fixed_extr, C.c_float), _p(pb.K, C.c_float), Np, _p(pt, C.c_double), No,
                            _p(pb.obs_cam, C.c_int32), _p(pb.obs_pt, C.c_int32), _p(pb.obs_uv, C.c_float),
                            C.byref(cost), _p(res, C.c_double), _p(Jc, C.c_double), _p(Jp, C.c_double),
                            _p(gc, C.c_double), _p(gp, C.c_double))
    if rc:
        raise RuntimeError("orc_evaluate failed")
    return dict(cost=cost.value, residuals=res, jac_cam=Jc, jac_pt=Jp, grad_cam=gc, grad_pt=gp)


def solve(pb, opt=None, iters_cap=None):
    """ceres::Solve restatement.  Returns (cam, pt, summary, iterations[list of dict])."""
    opt = opt or default_options()
    Nc, Np, No = _prob_args(pb)
    cam, pt = pb.cam.copy(), pb.pt.copy()
    cap = iters_cap or (opt.max_num_iterations + 2)
    its = (Iteration * cap)()
    sm = Summary()
    rc = lib().orc_solve(C.byref(opt), Nc, _p(cam, C.c_double), _p(pb.cam_fixed, C.c_uint8),
                         _p(pb.fixed_extr, C.c_float), _p(pb.K, C.c_float), Np, _p(pt, C.c_double), No,
                         _p(pb.obs_cam, C.c_int32), _p(pb.obs_pt, C.c_int32), _p(pb.obs_uv, C.c_float),
                         C.byref(sm), its, cap)
    if rc:
        raise RuntimeError("orc_solve failed")
    return cam, pt, sm, [_it_dict(its[i]) for i in range(min(sm.num_iterations, cap))]


def _it_dict(it):
    return {f: getattr(it, f) for f, _ in Iteration._fields_}


def schur_system(pb, radius, pt_begin, pt_end, opt=None):
    opt = opt or default_options()
    Nc, Np, No = _prob_args(pb)
    S, rhs = np.zeros((6 * Nc, 6 * Nc)), np.zeros(6 * Nc)
    cost = C.c_double(0)
    f = lib().orc_schur_system
    f.argtypes = [C.c_void_p, C.c_double] + [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                              C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                              C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    rc = f(C.addressof(opt), float(radius), Nc, pb.cam.ctypes.data, pb.cam_fixed.ctypes.data,
           pb.fixed_extr.ctypes.data, pb.K.ctypes.data, Np, pb.pt.ctypes.data, No, pb.obs_cam.ctypes.data,
           pb.obs_pt.ctypes.data, pb.obs_uv.ctypes.data, int(pt_begin), int(pt_end), S.ctypes.data,
           rhs.ctypes.data, C.addressof(cost))
    if rc:
        raise RuntimeError("orc_schur_system failed")
    return S, rhs, cost.value


def solve_pose_only_batch(batch, opt=None, mask=None):
    opt = opt or default_options(max_num_iterations=20)
    B = batch.cam.shape[0]
    cam = batch.cam.copy()
    cap = opt.max_num_iterations + 2
    its = (Iteration * (cap * B))()
    sms = (Summary * B)()
    m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    rc = lib().orc_solve_pose_only_batch(C.byref(opt), B, _p(cam, C.c_double), _p(batch.K, C.c_float),
                                         _p(batch.prob_ptr, C.c_int32), _p(batch.Xh, C.c_float),
                                         _p(batch.uv, C.c_float), _p(m, C.c_uint8), sms, its, cap)
    if rc:
        raise RuntimeError("orc_solve_pose_only_batch failed")
    iters = [[_it_dict(its[b * cap + i]) for i in range(min(sms[b].num_iterations, cap))] for b in range(B)]
    return cam, list(sms), iters


def prune_frame(pose, K, kp_xy, kp_octave, pt_xyz, pt_min, pt_max, has_point, outlier, consider_outlier=True):
    pose = np.ascontiguousarray(pose, np.float32).reshape(16); K = np.ascontiguousarray(K, np.float32).reshape(9)
    kp_xy = np.ascontiguousarray(kp_xy, np.float32); kp_octave = np.ascontiguousarray(kp_octave, np.int32)
    pt_xyz = np.ascontiguousarray(pt_xyz, np.float32)
    pt_min = np.ascontiguousarray(pt_min, np.float32); pt_max = np.ascontiguousarray(pt_max, np.float32)
    has_point = np.ascontiguousarray(has_point, np.uint8)
    out = np.ascontiguousarray(outlier, np.uint8).copy()
    lib().orc_prune_frame(_p(pose, C.c_float), _p(K, C.c_float), kp_xy.shape[0], _p(kp_xy, C.c_float),
                          _p(kp_octave, C.c_int32), _p(pt_xyz, C.c_float), _p(pt_min, C.c_float),
                          _p(pt_max, C.c_float), _p(has_point, C.c_uint8), _p(out, C.c_uint8), int(consider_outlier))
    return out


def pose_to_cam(pose):
    pose = np.ascontiguousarray(pose, np.float32).reshape(16)
    cam, extr = np.zeros(6), np.zeros(16, np.float32)
    lib().orc_pose_to_cam(_p(pose, C.c_float), _p(cam, C.c_double), _p(extr, C.c_float))
    return cam, extr.reshape(4, 4)


def cam_to_pose(cam):
    cam = np.ascontiguousarray(cam, np.float64)
    pose = np.zeros(16, np.float32)
    lib().orc_cam_to_pose(_p(cam, C.c_double), _p(pose, C.c_float))
    return pose.reshape(4, 4)


def cholesky_solve(A, b, num_threads=4):
    A = np.ascontiguousarray(A, np.float64).copy(); b = np.ascontiguousarray(b, np.float64).copy()
    rc = lib().orc_cholesky_solve(A.shape[0], _p(A, C.c_double), _p(b, C.c_double), num_threads)
    return rc, b
