// ProblemIO.h -- on-disk formats either side of the BA hot path:
//   * BAL text problems ("bundle adjustment in the large": `cams pts obs` header, one observation per line, then
//     9 numbers per camera and 3 per point) -> the flat arrays ba_set_problem takes (include/ba_b200.h);
//   * a flat binary dump of exactly those arrays (checkpoint / test-fixture format; written after a solve it is
//     a resumable state: parameters + structure);
//   * the trajectory file the reference writes at the end of a run (ba_project/src/ba/BundleAdjustment.cpp:248-268):
//     one line per keyframe, "timestamp tx ty tz qx qy qz qw", fixed notation with 4 decimals, pose = camera->world.
// The reference has no BAL reader; its only on-disk output on this path is the trajectory file.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

struct ba_handle;

namespace ba_io {

/** The input of ba_set_problem, owned by the host. */
struct FlatProblem {
    std::vector<double> cam;            // 6*Nc angle-axis + translation (world->camera)
    std::vector<uint8_t> cam_fixed;     // Nc
    std::vector<float> fixed_extr;      // 16*Nc row-major world->camera (read where cam_fixed)
    std::vector<float> K;               // 9*Nc row-major
    std::vector<double> pt;             // 3*Np
    std::vector<int32_t> obs_cam, obs_pt;
    std::vector<float> obs_uv;          // 2*No
    std::vector<double> bal_distortion; // 2*Nc (k1, k2) as read from a BAL file; carried, never applied (see readBAL)

    int32_t numCameras() const { return (int32_t)cam_fixed.size(); }
    int32_t numPoints() const { return (int32_t)(pt.size() / 3); }
    int32_t numObservations() const { return (int32_t)obs_cam.size(); }
    /** Sizes consistent and every index in range. */
    bool valid(std::string *why = nullptr) const;
};

/** BAL text -> FlatProblem.  BAL's camera is P = R(w) X + t, p = -P / P.z, pixel = f * r(p) * p with
 *  r(p) = 1 + k1 |p|^2 + k2 |p|^4.  The reference's camera model has no distortion (Optimizer.h:64-74: K * (R X + t)
 *  hnormalised), so the pinhole part maps EXACTLY onto a per-camera K = diag(-f, -f, 1) and k1, k2 are carried in
 *  bal_distortion without being applied; `n_distorted` (optional) receives the number of cameras with k1 or k2 != 0.
 *  fixFirstCamera: camera 0 becomes the gauge anchor the way frame ID 0 is in the reference (Optimizer.cpp:315-321):
 *  its observations turn into point-only residuals with float32 extrinsics. */
bool readBAL(const std::string &path, FlatProblem &out, std::string *err = nullptr, bool fixFirstCamera = true,
             int *n_distorted = nullptr);
/** FlatProblem -> BAL text (17 significant digits: doubles round-trip bit-exactly).  Needs K of the form
 *  [[f,0,cx],[0,f,cy],[0,0,1]] per camera (f of either sign); pixels are written relative to the principal point in
 *  BAL's sign convention.  Fixed cameras are written with the angle-axis pose in `cam`. */
bool writeBAL(const std::string &path, const FlatProblem &p, std::string *err = nullptr);

/** Flat binary dump, little-endian: 8-byte magic "BAB200P\1", int32 {Nc, Np, No, has_distortion}, then the arrays in
 *  the order of FlatProblem's members, each padded to 8 bytes. */
bool writeBinary(const std::string &path, const FlatProblem &p, std::string *err = nullptr);
bool readBinary(const std::string &path, FlatProblem &out, std::string *err = nullptr);

/** ba_set_problem on the arrays of `p` (status code of include/ba_b200.h). */
int setProblem(ba_handle *h, const FlatProblem &p);
/** Checkpoint after a solve: ba_get_cameras / ba_get_points into `p`.  Fixed cameras keep their entries. */
int fetchParameters(ba_handle *h, FlatProblem &p);

/** Rotation matrix -> quaternion (x, y, z, w), float32, the branch structure of Eigen::Quaternionf(Matrix3f) which the
 *  reference's writer uses (BundleAdjustment.cpp:261).  R row-major. */
void rotationToQuaternion(const float R[9], float q[4]);

/** The reference's trajectory writer on flat arrays: pose = camera->world 4x4 float row-major per keyframe, in keyframe
 *  order.  Writes `<outputPath>_estimatedPoses.txt` exactly like BundleAdjustment.cpp:251-264. */
bool writeTrajectoryTUM(const std::string &outputPath, int n, const double *timestamps, const float *poses,
                        std::string *err = nullptr);

/** Same, straight from a SceneMap (the reference's or the mock: getKeyFrames / getKeyFrameCount / getPose /
 *  getTimeStamp). */
template <class SceneMapT>
bool writeTrajectoryTUM(const std::string &outputPath, SceneMapT *map, std::string *err = nullptr) {
    auto frames = map->getKeyFrames();
    const int n = map->getKeyFrameCount();
    std::vector<double> ts((size_t)n);
    std::vector<float> poses(16 * (size_t)n);
    for (int i = 0; i < n; ++i) {
        auto pose = frames[i]->getPose();
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) poses[16 * (size_t)i + 4 * r + c] = pose(r, c);
        ts[i] = frames[i]->getTimeStamp();
    }
    return writeTrajectoryTUM(outputPath, n, ts.data(), poses.data(), err);
}

}  // namespace ba_io
