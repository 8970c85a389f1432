"""ctypes binding of libba_b200.so (include/ba_b200.h).  Python here is plumbing only: numpy arrays in, numpy
arrays out; every compute call goes through the C ABI into the CUDA kernels.  There is no CPU fallback: if the
shared library is missing, import of this module raises."""
from __future__ import annotations

import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libba_b200.so")

DENSE_SCHUR, ITERATIVE_SCHUR = 0, 1
CONVERGENCE, NO_CONVERGENCE, FAILURE = 0, 1, 2

# every symbol include/ba_b200.h declares (checked by tests/test_abi.py against the header text)
EXPORTS = [
    "ba_version", "ba_last_error", "ba_default_options", "ba_create", "ba_destroy", "ba_set_options",
    "ba_set_problem", "ba_set_parameters", "ba_reset_parameters", "ba_solve", "ba_get_cameras", "ba_get_points",
    "ba_get_iterations", "ba_evaluate", "ba_evaluate_device", "ba_schur_system", "ba_solve_pose_only_batch",
    "ba_prune_frames", "ba_prune_frames_noalloc", "ba_comm_unique_id", "ba_comm_init", "ba_get_launch_count", "ba_set_profiling", "ba_dense_spd_solve",
    "ba_get_reduced_system_stats", "ba_tile_elimination_order", "ba_get_kernel_profile", "ba_set_reduced_system_mode",
    "ba_solve_pose_only_outer", "ba_solve_outer", "ba_set_reduced_system_tile", "ba_dense_spd_solve_tile", "ba_set_pcg_mode", "ba_release_workspace", "ba_workspace_launch_count",
    "ba_partition_landmarks", "ba_inverse4x4_f32", "ba_angle_axis_to_rotation", "ba_rotation_to_angle_axis",
    "ba_pose_to_cam", "ba_cam_to_pose",
]


class Options(C.Structure):
    _fields_ = [
        ("max_num_iterations", C.c_int32), ("num_threads", C.c_int32), ("linear_solver", C.c_int32),
        ("jacobi_scaling", C.c_int32), ("max_num_consecutive_invalid_steps", C.c_int32),
        ("max_linear_solver_iterations", C.c_int32), ("use_huber", C.c_int32), ("pad0", C.c_int32),
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


class BAError(RuntimeError):
    pass


if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(nvcc, sm_100a).  bundleadjustment_b200 has no CPU fallback.")

lib = C.CDLL(LIB_PATH)
lib.ba_last_error.restype = C.c_char_p
lib.ba_tile_elimination_order.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
lib.ba_get_kernel_profile.argtypes = [C.c_void_p, C.c_void_p]
lib.ba_get_reduced_system_stats.argtypes = [C.c_void_p, C.c_void_p]
lib.ba_dense_spd_solve.argtypes = [C.c_int, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]
lib.ba_dense_spd_solve_tile.argtypes = [C.c_int, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]
lib.ba_prune_frames_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_int32, C.c_int32] + [C.c_void_p] * 10 + [C.c_int32]
lib.ba_schur_system.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]


def _check(rc: int, what: str):
    if rc != 0:
        raise BAError(f"{what} failed (code {rc}): {lib.ba_last_error().decode()}")


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def default_options(**kw) -> Options:
    o = Options()
    lib.ba_default_options(C.byref(o))
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


def it_dict(it) -> dict:
    return {f: getattr(it, f) for f, _ in Iteration._fields_}


class Solver:
    """One ceres::Problem + ceres::Solve replacement on one GPU (thin wrapper over a ba_handle)."""

    def __init__(self, options: Options | None = None, device: int = 0):
        self.options = options or default_options()
        self._h = C.c_void_p()
        _check(lib.ba_create(C.byref(self.options), int(device), C.byref(self._h)), "ba_create")
        self.shape = None

    def close(self):
        if self._h:
            lib.ba_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_options(self, options: Options):
        self.options = options
        _check(lib.ba_set_options(self._h, C.byref(options)), "ba_set_options")

    def comm_init(self, n_ranks: int, rank: int, unique_id: bytes):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        _check(lib.ba_comm_init(self._h, n_ranks, rank, buf), "ba_comm_init")

    def set_problem(self, pb):
        cam = np.ascontiguousarray(pb.cam, np.float64)
        fixed = np.ascontiguousarray(pb.cam_fixed, np.uint8)
        extr = np.ascontiguousarray(pb.fixed_extr, np.float32)
        K = np.ascontiguousarray(pb.K, np.float32)
        pt = np.ascontiguousarray(pb.pt, np.float64)
        oc = np.ascontiguousarray(pb.obs_cam, np.int32)
        op = np.ascontiguousarray(pb.obs_pt, np.int32)
        uv = np.ascontiguousarray(pb.obs_uv, np.float32)
        self.shape = (cam.shape[0], pt.shape[0], oc.shape[0])
        _check(lib.ba_set_problem(self._h, self.shape[0], _ptr(cam), _ptr(fixed), _ptr(extr), _ptr(K), self.shape[1],
                                  _ptr(pt), self.shape[2], _ptr(oc), _ptr(op), _ptr(uv)), "ba_set_problem")

    def set_parameters(self, cam, pt):
        cam = np.ascontiguousarray(cam, np.float64)
        pt = np.ascontiguousarray(pt, np.float64)
        _check(lib.ba_set_parameters(self._h, _ptr(cam), _ptr(pt)), "ba_set_parameters")

    def reset_parameters(self):
        _check(lib.ba_reset_parameters(self._h), "ba_reset_parameters")

    def solve(self) -> Summary:
        sm = Summary()
        _check(lib.ba_solve(self._h, C.byref(sm)), "ba_solve")
        return sm

    def set_reduced_system_mode(self, natural_order: bool = False, dense_tiles: bool = False):
        """Explicit (never environment-driven) switches of the reduced-system factorisation; call before set_problem."""
        _check(lib.ba_set_reduced_system_mode(self._h, int(natural_order), int(dense_tiles)), "ba_set_reduced_system_mode")

    def set_reduced_system_tile(self, tile_edge: int = 0):
        """Cholesky tile edge of the reduced system: 0 = automatic, 64 or 128 = forced."""
        _check(lib.ba_set_reduced_system_tile(self._h, int(tile_edge)), "ba_set_reduced_system_tile")

    def set_pcg_mode(self, mode: int = 0):
        """PCG matvec accumulation: 0 automatic, 1 ordered (bit-reproducible), 2 f64 atomics."""
        _check(lib.ba_set_pcg_mode(self._h, int(mode)), "ba_set_pcg_mode")

    def set_profiling(self, on: bool = True):
        _check(lib.ba_set_profiling(self._h, int(on)), "ba_set_profiling")

    def launch_count(self) -> int:
        return int(lib.ba_get_launch_count(self._h))

    def kernel_profile(self) -> dict:
        out = np.zeros(4)
        _check(lib.ba_get_kernel_profile(self._h, _ptr(out)), "ba_get_kernel_profile")
        return {"cholesky_seconds": float(out[0]), "cholesky_launches": int(out[1]),      # dense Schur: the Cholesky kernel
                "seconds": float(out[0]), "launches": int(out[1])}                        # (PCG: the matvec launches that did work)

    def reduced_system_stats(self) -> dict:
        """Order of the reduced camera system, filled / total 128x128 tiles of its Cholesky factor, flops per solve."""
        out = np.zeros(4)
        _check(lib.ba_get_reduced_system_stats(self._h, _ptr(out)), "ba_get_reduced_system_stats")
        return {"n": int(out[0]), "filled_tiles": int(out[1]), "lower_tiles": int(out[2]), "flops": float(out[3])}

    def cameras(self, out: np.ndarray | None = None) -> np.ndarray:
        """out (optional): a C-contiguous float64 [Nc, 6] destination, e.g. page-locked memory."""
        if out is None:
            out = np.zeros((self.shape[0], 6))
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.shape == (self.shape[0], 6)
        _check(lib.ba_get_cameras(self._h, _ptr(out)), "ba_get_cameras")
        return out

    def points(self, out: np.ndarray | None = None) -> np.ndarray:
        if out is None:
            out = np.zeros((self.shape[1], 3))
        assert out.dtype == np.float64 and out.flags.c_contiguous and out.shape == (self.shape[1], 3)
        _check(lib.ba_get_points(self._h, _ptr(out)), "ba_get_points")
        return out

    def iterations(self) -> list[dict]:
        cap = self.options.max_num_iterations + 2
        its = (Iteration * cap)()
        n = C.c_int32(0)
        _check(lib.ba_get_iterations(self._h, its, cap, C.byref(n)), "ba_get_iterations")
        return [it_dict(its[i]) for i in range(n.value)]

    def evaluate(self, jacobians: bool = True, gradient: bool = True) -> dict:
        Nc, Np, No = self.shape
        cost = C.c_double(0)
        res = np.zeros((No, 2))
        Jc = np.zeros((No, 2, 6)) if jacobians else None
        Jp = np.zeros((No, 2, 3)) if jacobians else None
        gc = np.zeros((Nc, 6)) if gradient else None
        gp = np.zeros((Np, 3)) if gradient else None
        _check(lib.ba_evaluate(self._h, C.byref(cost), _ptr(res), _ptr(Jc), _ptr(Jp), _ptr(gc), _ptr(gp)), "ba_evaluate")
        return dict(cost=cost.value, residuals=res, jac_cam=Jc, jac_pt=Jp, grad_cam=gc, grad_pt=gp)

    def evaluate_device_ms(self, repeats: int = 10) -> float:
        ms = C.c_float(0)
        _check(lib.ba_evaluate_device(self._h, repeats, C.byref(ms)), "ba_evaluate_device")
        return float(ms.value)

    def schur_system(self, radius: float):
        n = 6 * self.shape[0]
        S, rhs = np.zeros((n, n)), np.zeros(n)
        cost = C.c_double(0)
        _check(lib.ba_schur_system(self._h, float(radius), _ptr(S), _ptr(rhs), C.addressof(cost)), "ba_schur_system")
        return S, rhs, cost.value


def solve_pose_only_batch(batch, options: Options | None = None, mask=None, device: int = 0, want_iterations=True):
    """MotionOnlyBAOptimizerAngles' ceres::Solve for B frames in one launch."""
    opt = options or default_options(max_num_iterations=20)
    B = batch.cam.shape[0]
    cam = np.ascontiguousarray(batch.cam, np.float64).copy()
    K = np.ascontiguousarray(batch.K, np.float32)
    ptr = np.ascontiguousarray(batch.prob_ptr, np.int32)
    Xh = np.ascontiguousarray(batch.Xh, np.float32)
    uv = np.ascontiguousarray(batch.uv, np.float32)
    m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    cap = opt.max_num_iterations + 2
    its = (Iteration * (cap * B))() if want_iterations else None
    sms = (Summary * B)()
    _check(lib.ba_solve_pose_only_batch(C.byref(opt), device, B, _ptr(cam), _ptr(K), _ptr(ptr), _ptr(Xh), _ptr(uv),
                                        _ptr(m), sms, its, cap), "ba_solve_pose_only_batch")
    iters = None
    if want_iterations:
        iters = [[it_dict(its[b * cap + i]) for i in range(min(sms[b].num_iterations, cap))] for b in range(B)]
    return cam, list(sms), iters


def prune_frames(pose, K, frame_ptr, kp_xy, kp_octave, pt_xyz, pt_min, pt_max, has_point, outlier,
                 consider_outlier=True, device: int = 0):
    pose = np.ascontiguousarray(pose, np.float32); K = np.ascontiguousarray(K, np.float32)
    frame_ptr = np.ascontiguousarray(frame_ptr, np.int32)
    kp_xy = np.ascontiguousarray(kp_xy, np.float32); kp_octave = np.ascontiguousarray(kp_octave, np.int32)
    pt_xyz = np.ascontiguousarray(pt_xyz, np.float32)
    pt_min = np.ascontiguousarray(pt_min, np.float32); pt_max = np.ascontiguousarray(pt_max, np.float32)
    has_point = np.ascontiguousarray(has_point, np.uint8)
    out = np.ascontiguousarray(outlier, np.uint8).copy()
    _check(lib.ba_prune_frames(device, frame_ptr.shape[0] - 1, _ptr(pose), _ptr(K), _ptr(frame_ptr), _ptr(kp_xy),
                               _ptr(kp_octave), _ptr(pt_xyz), _ptr(pt_min), _ptr(pt_max), _ptr(has_point), _ptr(out),
                               int(consider_outlier)), "ba_prune_frames")
    return out


PRUNE_NOALLOC_MAX_FRAMES = 4096          # BA_PRUNE_NOALLOC_MAX_FRAMES (include/ba_b200.h)


def prune_frames_noalloc(n_frames: int, n_kp: int, d_pose: int, d_K: int, d_frame_ptr: int, d_kp_xy: int, d_kp_octave: int,
                         d_pt_xyz: int, d_pt_min: int, d_pt_max: int, d_has_point: int, d_outlier: int, consider_outlier=True,
                         device: int = 0, stream: int = 0) -> None:
    """ba_prune_frames_noalloc: every d_* is a DEVICE address (int), `stream` a cudaStream_t handle (int, 0 = default stream).
    Enqueues only -- synchronise the stream before reading d_outlier.  Allocates nothing (static device scratch)."""
    _check(lib.ba_prune_frames_noalloc(int(device), C.c_void_p(stream or None), int(n_frames), int(n_kp), *[C.c_void_p(int(a)) for a in (
        d_pose, d_K, d_frame_ptr, d_kp_xy, d_kp_octave, d_pt_xyz, d_pt_min, d_pt_max, d_has_point, d_outlier)], int(consider_outlier)),
        "ba_prune_frames_noalloc")


def solve_pose_only_outer(pose, K, frame_ptr, kp_xy, kp_octave, pt_xyz, pt_min, pt_max, has_point, outlier, n_outer=4,
                          options: Options | None = None, device: int = 0, pose_out=None, outlier_out=None):
    """ba_solve_pose_only_outer: returns (pose [F,4,4] f32, outlier u8, summaries of the last outer iteration, device ms,
    LM iterations of all frames and outer iterations).
    pose_out / outlier_out (optional): caller-owned (e.g. page-locked) result buffers."""
    F = int(np.asarray(frame_ptr).shape[0]) - 1
    P = np.ascontiguousarray(pose, np.float32).reshape(F, 16)
    if pose_out is None:
        pose_out = P.copy()
    else:
        pose_out.reshape(F, 16)[...] = P
    o = np.ascontiguousarray(outlier, np.uint8)
    if outlier_out is None:
        outlier_out = o.copy()
    else:
        outlier_out[...] = o
    args = [np.ascontiguousarray(K, np.float32), np.ascontiguousarray(frame_ptr, np.int32), np.ascontiguousarray(kp_xy, np.float32),
            np.ascontiguousarray(kp_octave, np.int32), np.ascontiguousarray(pt_xyz, np.float32), np.ascontiguousarray(pt_min, np.float32),
            np.ascontiguousarray(pt_max, np.float32), np.ascontiguousarray(has_point, np.uint8)]
    sms = (Summary * max(F, 1))()
    ms = C.c_float(0)
    lm = C.c_int64(0)
    opt = options or default_options(max_num_iterations=20)
    _check(lib.ba_solve_pose_only_outer(C.byref(opt), int(device), F, _ptr(pose_out), *[_ptr(a) for a in args], _ptr(outlier_out),
                                        int(n_outer), sms, C.byref(ms), C.byref(lm)), "ba_solve_pose_only_outer")
    return pose_out.reshape(F, 4, 4), outlier_out, list(sms)[:F], float(ms.value), int(lm.value)


def workspace_launch_count(device: int = 0) -> int:
    return int(lib.ba_workspace_launch_count(int(device)))


def dense_spd_solve(A, b, device: int = 0, repeats: int = 1, tile_edge: int = 0):
    """tile_edge: 0 = automatic (by the tile sparsity of A), 64 / 128 = that instantiation of the Cholesky kernel"""
    A = np.ascontiguousarray(A, np.float64); b = np.ascontiguousarray(b, np.float64)
    x = np.zeros_like(b)
    ms = C.c_float(0)
    _check(lib.ba_dense_spd_solve_tile(device, A.shape[0], _ptr(A), _ptr(b), _ptr(x), C.addressof(ms), repeats, int(tile_edge)), "ba_dense_spd_solve")
    return x, float(ms.value)


# ---- host-only helpers -------------------------------------------------------------------------------------
def tile_elimination_order(adj):
    """Nested-dissection order of the Cholesky tiles + {depth, filled tiles} before/after (host only)."""
    adj = np.ascontiguousarray(adj, np.uint8)
    G = adj.shape[0]
    order = np.zeros(G, np.int32)
    stats = np.zeros(4)
    _check(lib.ba_tile_elimination_order(G, _ptr(adj), _ptr(order), _ptr(stats)), "ba_tile_elimination_order")
    return order, {"depth_natural": int(stats[0]), "depth_ordered": int(stats[1]),
                   "filled_natural": int(stats[2]), "filled_ordered": int(stats[3])}


def comm_unique_id() -> bytes:
    buf = (C.c_uint8 * 128)()
    _check(lib.ba_comm_unique_id(buf), "ba_comm_unique_id")
    return bytes(buf)


def partition_landmarks(n_pt: int, obs_pt, n_ranks: int, quadratic: bool = False) -> np.ndarray:
    obs_pt = np.ascontiguousarray(obs_pt, np.int32)
    cut = np.zeros(n_ranks + 1, np.int32)
    _check(lib.ba_partition_landmarks(n_pt, obs_pt.shape[0], _ptr(obs_pt), n_ranks, int(quadratic), _ptr(cut)),
           "ba_partition_landmarks")
    return cut


def pose_to_cam(pose):
    pose = np.ascontiguousarray(pose, np.float32).reshape(16)
    cam, extr = np.zeros(6), np.zeros(16, np.float32)
    lib.ba_pose_to_cam(_ptr(pose), _ptr(cam), _ptr(extr))
    return cam, extr.reshape(4, 4)


def cam_to_pose(cam):
    cam = np.ascontiguousarray(cam, np.float64)
    pose = np.zeros(16, np.float32)
    lib.ba_cam_to_pose(_ptr(cam), _ptr(pose))
    return pose.reshape(4, 4)


def inverse4x4_f32(M):
    M = np.ascontiguousarray(M, np.float32).reshape(16)
    out = np.zeros(16, np.float32)
    lib.ba_inverse4x4_f32(_ptr(M), _ptr(out))
    return out.reshape(4, 4)
