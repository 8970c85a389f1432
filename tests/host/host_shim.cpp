// host_shim.cpp -- TEST INFRASTRUCTURE: a C-callable driver that builds the mock SceneMap / Frame / MapPoint object
// graph (bundleadjustment_b200/host/ba_model_mock.h) from flat arrays, runs the drop-in optimizer classes
// (bundleadjustment_b200/host/Optimizer.h) exactly the way the reference's callers do (BundleAdjustment.cpp:196-207,
// 228-230; SfMHelper.cpp:106-117) and flattens the result back so Python tests can compare it with the oracle flow.
#include "../../bundleadjustment_b200/host/Optimizer.h"

#include <cstdint>
#include <unordered_map>
#include <cstring>

namespace {
struct Graph {
    std::vector<std::shared_ptr<Frame>> frames;
    std::vector<std::shared_ptr<MapPoint>> points;
    SceneMap map;
};

void build(Graph &g, int Nc, const float *pose, const float *K, const int32_t *frame_id, const int32_t *kp_ptr, const float *kp_xy,
           const int32_t *kp_octave, const int32_t *kp_point, const uint8_t *kp_outlier, int Np, const float *pt_xyz,
           const float *pt_min, const float *pt_max) {
    for (int f = 0; f < Nc; ++f) {
        Matrix4f P; Matrix3f Km;
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) P(r, c) = pose[16 * f + 4 * r + c];
        for (int r = 0; r < 3; ++r) for (int c = 0; c < 3; ++c) Km(r, c) = K[9 * f + 3 * r + c];
        auto fr = std::make_shared<Frame>(P, Km, frame_id[f]);
        std::vector<cv::KeyPoint> kps(kp_ptr[f + 1] - kp_ptr[f]);
        for (size_t i = 0; i < kps.size(); ++i) {
            const int k = kp_ptr[f] + (int)i;
            kps[i].pt.x = kp_xy[2 * k]; kps[i].pt.y = kp_xy[2 * k + 1]; kps[i].octave = kp_octave[k];
        }
        fr->setKeypoints(kps);
        g.frames.push_back(fr);
    }
    g.points.assign(Np, nullptr);
    for (int f = 0; f < Nc; ++f)
        for (int k = kp_ptr[f]; k < kp_ptr[f + 1]; ++k) {
            const int i = k - kp_ptr[f];
            if (kp_outlier[k]) g.frames[f]->setOutlier(i);
            const int p = kp_point[k];
            if (p < 0) continue;
            if (!g.points[p]) {
                Vector3f X; X(0) = pt_xyz[3 * p]; X(1) = pt_xyz[3 * p + 1]; X(2) = pt_xyz[3 * p + 2];
                g.points[p] = std::make_shared<MapPoint>(X, g.frames[f], (size_t)i);
                g.points[p]->setDistanceBounds(pt_min[p], pt_max[p]);
            } else {
                g.points[p]->addObservation(g.frames[f], (size_t)i);
            }
            g.frames[f]->addAssociatedMapPoint(i, g.points[p]);
        }
    for (int p = 0; p < Np; ++p)
        if (!g.points[p]) {     // unobserved map point (legal: null observing frame, MapPoint.cpp:11)
            Vector3f X; X(0) = pt_xyz[3 * p]; X(1) = pt_xyz[3 * p + 1]; X(2) = pt_xyz[3 * p + 2];
            g.points[p] = std::make_shared<MapPoint>(X, nullptr, 0);
            g.points[p]->setDistanceBounds(pt_min[p], pt_max[p]);
        }
    for (auto &f : g.frames) g.map.addKeyFrame(f);
    for (auto &p : g.points) g.map.addMapPoint(p);
}

void dump(Graph &g, const int32_t *kp_ptr, float *pose_out, float *pt_out, uint8_t *outlier_out, int32_t *kp_point_out, uint8_t *pt_alive) {
    std::unordered_map<const MapPoint *, int> idx;
    for (size_t p = 0; p < g.points.size(); ++p) idx[g.points[p].get()] = (int)p;
    for (size_t f = 0; f < g.frames.size(); ++f) {
        const Matrix4f P = g.frames[f]->getPose();
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) pose_out[16 * f + 4 * r + c] = P(r, c);
        for (int k = kp_ptr[f]; k < kp_ptr[f + 1]; ++k) {
            outlier_out[k] = g.frames[f]->isOutlier(k - kp_ptr[f]) ? 1 : 0;
            auto mp = g.frames[f]->getMapPoint(k - kp_ptr[f]);
            kp_point_out[k] = mp ? idx[mp.get()] : -1;
        }
    }
    std::memset(pt_alive, 0, g.points.size());
    for (auto &mp : g.map.getMapPoints()) pt_alive[idx[mp.get()]] = 1;
    for (size_t p = 0; p < g.points.size(); ++p) {
        const Vector3f X = g.points[p]->getPosition();
        pt_out[3 * p] = X(0); pt_out[3 * p + 1] = X(1); pt_out[3 * p + 2] = X(2);
    }
}
}  // namespace

#include <unordered_map>

// mode 0: GlobalBAOptimizerAngles   1: LocalBAOptimizerAngles (current = frame `current`, covisibility order `cov`)
// mode 2: MotionOnlyBAOptimizerAngles::optimizeCameraPose on every frame in turn   3: ::optimizeCameraPoses (batched)
extern "C" int ba_host_shim_run(int mode, int Nc, const float *pose, const float *K, const int32_t *frame_id, const int32_t *kp_ptr,
                                const float *kp_xy, const int32_t *kp_octave, const int32_t *kp_point, const uint8_t *kp_outlier,
                                int Np, const float *pt_xyz, const float *pt_min, const float *pt_max, int n_outer, int max_it,
                                int erase_outliers, int current, const int32_t *cov, int n_cov, float *pose_out, float *pt_out,
                                uint8_t *outlier_out, int32_t *kp_point_out, uint8_t *pt_alive, double *info /*[4]*/) {
    Graph g;
    build(g, Nc, pose, K, frame_id, kp_ptr, kp_xy, kp_octave, kp_point, kp_outlier, Np, pt_xyz, pt_min, pt_max);
    int status = 0;
    if (mode == 0) {
        GlobalBAOptimizerAngles opt;
        opt.setNbOfIterations(n_outer); opt.setNbOfMaxItPerBA(max_it);
        opt.optimizeCamerasAndMapPoints(&g.map, erase_outliers != 0, current < 0 ? 0 : g.frames[current]->getID());
        status = opt.lastStatus();
        info[0] = opt.lastTerminationType(); info[1] = opt.lastInitialCost(); info[2] = opt.lastFinalCost(); info[3] = opt.lastNumIterations();
    } else if (mode == 1) {
        std::vector<std::shared_ptr<Frame>> order;
        for (int i = 0; i < n_cov; ++i) order.push_back(g.frames[cov[i]]);
        g.frames[current]->setOrderedCovisibilityFrames(order);
        LocalBAOptimizerAngles opt;
        opt.setNbOfIterations(n_outer); opt.setNbOfMaxItPerBA(max_it);
        opt.optimizeCamerasAndMapPoints(&g.map, g.frames[current], erase_outliers != 0);
        status = opt.lastStatus();
        info[0] = opt.lastTerminationType(); info[1] = opt.lastInitialCost(); info[2] = opt.lastFinalCost(); info[3] = opt.lastNumIterations();
    } else {
        MotionOnlyBAOptimizerAngles opt;       // SfMHelper::estimatePoseUsingBA sets 4 x 20 (SfMHelper.cpp:112-113)
        opt.setNbOfIterations(n_outer); opt.setNbOfMaxItPerBA(max_it);
        if (mode == 2) for (auto &f : g.frames) { opt.optimizeCameraPose(f); if (opt.lastStatus()) break; }
        else opt.optimizeCameraPoses(g.frames);
        status = opt.lastStatus();
        info[0] = opt.lastTerminationType(); info[1] = opt.lastInitialCost(); info[2] = opt.lastFinalCost(); info[3] = opt.lastNumIterations();
    }
    dump(g, kp_ptr, pose_out, pt_out, outlier_out, kp_point_out, pt_alive);
    return status;
}

// ---------------------------------------------------------------------------------------------------------------
// ProblemIO: format conversions, the trajectory writer through the SceneMap template, and the
// load -> solve -> checkpoint flow.  `err` receives a NUL-terminated message (cap bytes).
// ---------------------------------------------------------------------------------------------------------------
#include "../../bundleadjustment_b200/host/ProblemIO.h"
#include "../../include/ba_b200.h"

namespace {
int report(bool ok, const std::string &msg, char *err, int cap) {
    if (err && cap > 0) { std::strncpy(err, ok ? "" : msg.c_str(), (size_t)cap - 1); err[cap - 1] = '\0'; }
    return ok ? 0 : 1;
}
}  // namespace

// kind: 0 = BAL text, 1 = flat binary
extern "C" int ba_host_shim_convert(const char *src, int src_kind, const char *dst, int dst_kind, int fix_first, int32_t *shape /*[4]*/,
                                    char *err, int cap) {
    ba_io::FlatProblem p;
    std::string msg;
    int distorted = 0;
    bool ok = src_kind == 0 ? ba_io::readBAL(src, p, &msg, fix_first != 0, &distorted) : ba_io::readBinary(src, p, &msg);
    if (ok) ok = dst_kind == 0 ? ba_io::writeBAL(dst, p, &msg) : ba_io::writeBinary(dst, p, &msg);
    if (ok && shape) { shape[0] = p.numCameras(); shape[1] = p.numPoints(); shape[2] = p.numObservations(); shape[3] = distorted; }
    return report(ok, msg, err, cap);
}

extern "C" int ba_host_shim_trajectory(const char *prefix, int n, const int32_t *frame_id, const double *ts, const float *pose,
                                       char *err, int cap) {
    SceneMap map;
    for (int f = 0; f < n; ++f) {
        Matrix4f P;
        for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) P(r, c) = pose[16 * f + 4 * r + c];
        map.addKeyFrame(std::make_shared<Frame>(P, Matrix3f::Identity(), frame_id[f], ts[f]));
    }
    std::string msg;
    return report(ba_io::writeTrajectoryTUM(prefix, &map, &msg), msg, err, cap);
}

// readBinary -> ba_set_problem -> ba_solve -> fetch -> writeBinary (checkpoint).  info = {initial, final cost, iterations, termination}
extern "C" int ba_host_shim_solve_file(const char *src, const char *dst, int max_it, int linear_solver, double *info, char *err, int cap) {
    ba_io::FlatProblem p;
    std::string msg;
    if (!ba_io::readBinary(src, p, &msg)) return report(false, msg, err, cap);
    ba_options opt;
    ba_default_options(&opt);
    opt.max_num_iterations = max_it;
    opt.linear_solver = linear_solver;
    ba_handle *h = nullptr;
    ba_summary sm;
    int rc = ba_create(&opt, 0, &h);
    if (rc == BA_OK) rc = ba_io::setProblem(h, p);
    if (rc == BA_OK) rc = ba_solve(h, &sm);
    if (rc == BA_OK) rc = ba_io::fetchParameters(h, p);
    if (rc != BA_OK) msg = ba_last_error();
    if (h) ba_destroy(h);
    if (rc != BA_OK) return report(false, msg, err, cap);
    info[0] = sm.initial_cost; info[1] = sm.final_cost; info[2] = sm.num_iterations; info[3] = sm.termination_type;
    return report(ba_io::writeBinary(dst, p, &msg), msg, err, cap);
}
