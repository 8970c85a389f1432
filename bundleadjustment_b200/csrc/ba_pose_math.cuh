// ba_pose_math.cuh -- pose <-> parameter conversion of the reference's flattening / write-back, usable on the device.
//
// Reference lines: prepareConstraints (ba_project/src/ba/Optimizer.cpp:296-299, 462-465, 629-632: Matrix4f::inverse(),
// ceres::RotationMatrixToAngleAxis) and the write-back (Optimizer.cpp:262-267, 448-453, 590-595:
// ceres::AngleAxisToRotationMatrix, cast to float, Matrix4f::inverse()).  The host twins live in ba_host.cpp (compiled
// with -ffp-contract=off); the float32 inverse here uses explicit round-to-nearest intrinsics so that the device forms
// exactly the same float32 products and sums (no FMA contraction): pruneCorrespondences (Optimizer.cpp:6-79) is a
// float32 threshold test and must see the same extrinsics bit for bit.
#pragma once

#include "ba_device.cuh"

namespace ba {

__device__ __forceinline__ void dev_inverse4x4_f32(const float *__restrict__ m, float *__restrict__ out) {
#define M_(a, b) __fmul_rn(a, b)
#define S_(a, b) __fsub_rn(a, b)
#define A_(a, b) __fadd_rn(a, b)
    const float a00 = m[0], a01 = m[1], a02 = m[2], a03 = m[3];
    const float a10 = m[4], a11 = m[5], a12 = m[6], a13 = m[7];
    const float a20 = m[8], a21 = m[9], a22 = m[10], a23 = m[11];
    const float a30 = m[12], a31 = m[13], a32 = m[14], a33 = m[15];
    const float s0 = S_(M_(a00, a11), M_(a10, a01)), s1 = S_(M_(a00, a12), M_(a10, a02)), s2 = S_(M_(a00, a13), M_(a10, a03));
    const float s3 = S_(M_(a01, a12), M_(a11, a02)), s4 = S_(M_(a01, a13), M_(a11, a03)), s5 = S_(M_(a02, a13), M_(a12, a03));
    const float c5 = S_(M_(a22, a33), M_(a32, a23)), c4 = S_(M_(a21, a33), M_(a31, a23)), c3 = S_(M_(a21, a32), M_(a31, a22));
    const float c2 = S_(M_(a20, a33), M_(a30, a23)), c1 = S_(M_(a20, a32), M_(a30, a22)), c0 = S_(M_(a20, a31), M_(a30, a21));
    const float det = A_(S_(A_(A_(S_(M_(s0, c5), M_(s1, c4)), M_(s2, c3)), M_(s3, c2)), M_(s4, c1)), M_(s5, c0));
    const float id = __fdiv_rn(1.0f, det);
    // (x * p - y * q + z * r) * id, evaluated left to right like the host expression
#define T3(x, p, y, q, z, r) M_(A_(S_(M_(x, p), M_(y, q)), M_(z, r)), id)
    out[0] = T3(a11, c5, a12, c4, a13, c3);
    out[1] = M_(S_(A_(M_(-a01, c5), M_(a02, c4)), M_(a03, c3)), id);
    out[2] = T3(a31, s5, a32, s4, a33, s3);
    out[3] = M_(S_(A_(M_(-a21, s5), M_(a22, s4)), M_(a23, s3)), id);
    out[4] = M_(S_(A_(M_(-a10, c5), M_(a12, c2)), M_(a13, c1)), id);
    out[5] = T3(a00, c5, a02, c2, a03, c1);
    out[6] = M_(S_(A_(M_(-a30, s5), M_(a32, s2)), M_(a33, s1)), id);
    out[7] = T3(a20, s5, a22, s2, a23, s1);
    out[8] = T3(a10, c4, a11, c2, a13, c0);
    out[9] = M_(S_(A_(M_(-a00, c4), M_(a01, c2)), M_(a03, c0)), id);
    out[10] = T3(a30, s4, a31, s2, a33, s0);
    out[11] = M_(S_(A_(M_(-a20, s4), M_(a21, s2)), M_(a23, s0)), id);
    out[12] = M_(S_(A_(M_(-a10, c3), M_(a11, c1)), M_(a12, c0)), id);
    out[13] = T3(a00, c3, a01, c1, a02, c0);
    out[14] = M_(S_(A_(M_(-a30, s3), M_(a31, s1)), M_(a32, s0)), id);
    out[15] = T3(a20, s3, a21, s1, a22, s0);
#undef T3
#undef M_
#undef S_
#undef A_
}

// ceres::RotationMatrixToAngleAxis (via quaternion), column-major input
__device__ __forceinline__ void dev_rotation_to_angle_axis(const double *R, double *aa) {
#define RM(i, j) R[(i) + 3 * (j)]
    double q[4];
    const double trace = RM(0, 0) + RM(1, 1) + RM(2, 2);
    if (trace >= 0.0) {
        double t = sqrt(trace + 1.0);
        q[0] = 0.5 * t;
        t = 0.5 / t;
        q[1] = (RM(2, 1) - RM(1, 2)) * t;
        q[2] = (RM(0, 2) - RM(2, 0)) * t;
        q[3] = (RM(1, 0) - RM(0, 1)) * t;
    } else {
        int i = 0;
        if (RM(1, 1) > RM(0, 0)) i = 1;
        if (RM(2, 2) > RM(i, i)) i = 2;
        const int j = (i + 1) % 3, k = (j + 1) % 3;
        double t = sqrt(RM(i, i) - RM(j, j) - RM(k, k) + 1.0);
        q[i + 1] = 0.5 * t;
        t = 0.5 / t;
        q[0] = (RM(k, j) - RM(j, k)) * t;
        q[j + 1] = (RM(j, i) + RM(i, j)) * t;
        q[k + 1] = (RM(k, i) + RM(i, k)) * t;
    }
#undef RM
    const double sin2 = q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
    double k;
    if (sin2 > 0.0) {
        const double st = sqrt(sin2), ct = q[0];
        const double two_theta = 2.0 * ((ct < 0.0) ? atan2(-st, -ct) : atan2(st, ct));
        k = two_theta / st;
    } else {
        k = 2.0;
    }
    aa[0] = q[1] * k; aa[1] = q[2] * k; aa[2] = q[3] * k;
}

// Optimizer.cpp:296-299 / 462-465: extr = pose.inverse() (f32), rotation -> angle-axis (f64), translation -> f64
__device__ __forceinline__ void dev_pose_to_cam(const float *__restrict__ pose, double *__restrict__ cam, float *__restrict__ extr) {
    dev_inverse4x4_f32(pose, extr);
    double R[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) R[i + 3 * j] = (double)extr[i * 4 + j];
    dev_rotation_to_angle_axis(R, cam);
#pragma unroll
    for (int i = 0; i < 3; ++i) cam[3 + i] = (double)extr[i * 4 + 3];
}

// Optimizer.cpp:262-267 / 448-453: angle-axis -> R (f64) -> f32 extrinsics -> inverse -> Frame::setPose
__device__ __forceinline__ void dev_cam_to_pose(const double *__restrict__ cam, float *__restrict__ pose) {
    double R[9];
    rodrigues(cam, R, nullptr);
    float extr[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
#pragma unroll
        for (int j = 0; j < 3; ++j) extr[i * 4 + j] = (float)R[i + 3 * j];
        extr[i * 4 + 3] = (float)cam[3 + i];
    }
    dev_inverse4x4_f32(extr, pose);
}

} // namespace ba
