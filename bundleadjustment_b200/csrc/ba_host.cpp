// ba_host.cpp -- host-only parts of the C ABI: pose <-> parameter conversion used by the flattening / write-back
// of the drop-in optimizer classes, and the landmark partitioner for the multi-GPU mode.  No CUDA here.
//
// Reference lines: prepareConstraints' pose conversion (ba_project/src/ba/Optimizer.cpp:296-299, 462-465,
// 629-632: Matrix4f::inverse(), ceres::RotationMatrixToAngleAxis) and the write-back (Optimizer.cpp:262-267,
// 448-453, 590-595: ceres::AngleAxisToRotationMatrix, cast to float, Matrix4f::inverse()).
#include "../../include/ba_b200.h"

#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

extern "C" {

// General 4x4 inverse in float32 through the adjugate: 2x2 sub-determinants of the top and bottom row pairs
// (the formulation Eigen's compute_inverse_size4 uses), row-major in/out.
void ba_inverse4x4_f32(const float m[16], float out[16]) {
    const float a00 = m[0], a01 = m[1], a02 = m[2], a03 = m[3];
    const float a10 = m[4], a11 = m[5], a12 = m[6], a13 = m[7];
    const float a20 = m[8], a21 = m[9], a22 = m[10], a23 = m[11];
    const float a30 = m[12], a31 = m[13], a32 = m[14], a33 = m[15];
    const float s0 = a00 * a11 - a10 * a01, s1 = a00 * a12 - a10 * a02, s2 = a00 * a13 - a10 * a03;
    const float s3 = a01 * a12 - a11 * a02, s4 = a01 * a13 - a11 * a03, s5 = a02 * a13 - a12 * a03;
    const float c5 = a22 * a33 - a32 * a23, c4 = a21 * a33 - a31 * a23, c3 = a21 * a32 - a31 * a22;
    const float c2 = a20 * a33 - a30 * a23, c1 = a20 * a32 - a30 * a22, c0 = a20 * a31 - a30 * a21;
    const float det = s0 * c5 - s1 * c4 + s2 * c3 + s3 * c2 - s4 * c1 + s5 * c0;
    const float id = 1.0f / det;
    out[0] = (a11 * c5 - a12 * c4 + a13 * c3) * id;
    out[1] = (-a01 * c5 + a02 * c4 - a03 * c3) * id;
    out[2] = (a31 * s5 - a32 * s4 + a33 * s3) * id;
    out[3] = (-a21 * s5 + a22 * s4 - a23 * s3) * id;
    out[4] = (-a10 * c5 + a12 * c2 - a13 * c1) * id;
    out[5] = (a00 * c5 - a02 * c2 + a03 * c1) * id;
    out[6] = (-a30 * s5 + a32 * s2 - a33 * s1) * id;
    out[7] = (a20 * s5 - a22 * s2 + a23 * s1) * id;
    out[8] = (a10 * c4 - a11 * c2 + a13 * c0) * id;
    out[9] = (-a00 * c4 + a01 * c2 - a03 * c0) * id;
    out[10] = (a30 * s4 - a31 * s2 + a33 * s0) * id;
    out[11] = (-a20 * s4 + a21 * s2 - a23 * s0) * id;
    out[12] = (-a10 * c3 + a11 * c1 - a12 * c0) * id;
    out[13] = (a00 * c3 - a01 * c1 + a02 * c0) * id;
    out[14] = (-a30 * s3 + a31 * s1 - a32 * s0) * id;
    out[15] = (a20 * s3 - a21 * s1 + a22 * s0) * id;
}

// ceres::AngleAxisToRotationMatrix, column-major output
void ba_angle_axis_to_rotation(const double w[3], double R[9]) {
    const double theta2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
    if (theta2 > DBL_EPSILON) {
        const double theta = std::sqrt(theta2);
        const double wx = w[0] / theta, wy = w[1] / theta, wz = w[2] / theta;
        const double c = std::cos(theta), s = std::sin(theta);
        R[0] = c + wx * wx * (1.0 - c);       R[1] = wz * s + wx * wy * (1.0 - c);  R[2] = -wy * s + wx * wz * (1.0 - c);
        R[3] = wx * wy * (1.0 - c) - wz * s;  R[4] = c + wy * wy * (1.0 - c);       R[5] = wx * s + wy * wz * (1.0 - c);
        R[6] = wy * s + wx * wz * (1.0 - c);  R[7] = -wx * s + wy * wz * (1.0 - c); R[8] = c + wz * wz * (1.0 - c);
    } else {
        R[0] = 1.0;   R[1] = w[2];  R[2] = -w[1];
        R[3] = -w[2]; R[4] = 1.0;   R[5] = w[0];
        R[6] = w[1];  R[7] = -w[0]; R[8] = 1.0;
    }
}

// ceres::RotationMatrixToAngleAxis (via quaternion), column-major input
void ba_rotation_to_angle_axis(const double R[9], double aa[3]) {
#define RM(i, j) R[(i) + 3 * (j)]
    double q[4];
    const double trace = RM(0, 0) + RM(1, 1) + RM(2, 2);
    if (trace >= 0.0) {
        double t = std::sqrt(trace + 1.0);
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
        double t = std::sqrt(RM(i, i) - RM(j, j) - RM(k, k) + 1.0);
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
        const double st = std::sqrt(sin2), ct = q[0];
        const double two_theta = 2.0 * ((ct < 0.0) ? std::atan2(-st, -ct) : std::atan2(st, ct));
        k = two_theta / st;
    } else {
        k = 2.0;
    }
    aa[0] = q[1] * k; aa[1] = q[2] * k; aa[2] = q[3] * k;
}

void ba_pose_to_cam(const float pose[16], double cam[6], float extr_out[16]) {
    float extr[16];
    ba_inverse4x4_f32(pose, extr);
    double R[9];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) R[i + 3 * j] = (double)extr[i * 4 + j];
    ba_rotation_to_angle_axis(R, cam);
    for (int i = 0; i < 3; ++i) cam[3 + i] = (double)extr[i * 4 + 3];
    if (extr_out) std::memcpy(extr_out, extr, sizeof(extr));
}

void ba_cam_to_pose(const double cam[6], float pose_out[16]) {
    double R[9];
    ba_angle_axis_to_rotation(cam, R);
    float extr[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1};
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) extr[i * 4 + j] = (float)R[i + 3 * j];
        extr[i * 4 + 3] = (float)cam[3 + i];
    }
    ba_inverse4x4_f32(extr, pose_out);
}

int ba_partition_landmarks(int32_t n_pt, int32_t n_obs, const int32_t *obs_pt, int32_t n_ranks, int32_t quadratic, int32_t *cut) {
    if (n_pt < 0 || n_obs < 0 || n_ranks < 1 || !cut || (n_obs && !obs_pt)) return BA_ERR_INVALID;
    std::vector<long long> t(n_pt, 0);
    for (int i = 0; i < n_obs; ++i) {
        if (obs_pt[i] < 0 || obs_pt[i] >= n_pt) return BA_ERR_INVALID;
        if (i && obs_pt[i] < obs_pt[i - 1]) return BA_ERR_INVALID;      // must be landmark-sorted
        t[obs_pt[i]]++;
    }
    long double total = 0;
    std::vector<long double> w(n_pt);
    for (int p = 0; p < n_pt; ++p) {
        w[p] = quadratic ? (long double)t[p] * (t[p] + 1) / 2 : (long double)t[p];
        total += w[p];
    }
    cut[0] = 0;
    long double acc = 0;
    int r = 1;
    for (int p = 0; p < n_pt && r < n_ranks; ++p) {
        acc += w[p];
        while (r < n_ranks && acc >= total * r / n_ranks) cut[r++] = p + 1;
    }
    while (r <= n_ranks) cut[r++] = n_pt;
    cut[n_ranks] = n_pt;
    return BA_OK;
}

} // extern "C"
