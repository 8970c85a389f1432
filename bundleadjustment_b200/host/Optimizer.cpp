// Optimizer.cpp -- host adaptor: the reference's optimizer entry points on top of the C ABI (include/ba_b200.h).
//
// What stays on the host (it walks shared_ptr object graphs behind mutex-guarded getters, exactly like the reference):
// the flattening that prepareConstraints does (Optimizer.cpp:279-333, 459-498, 612-698), the write-back through
// Frame::setPose / MapPoint::setPosition (Optimizer.cpp:250-271, 446-455, 578-603), the outer iteration loop and the
// Local-BA window selection (Optimizer.cpp:502-546).  What moves to the GPU: everything ceres::Solve did, and
// pruneCorrespondences (one batched launch per outer iteration instead of a per-keypoint host loop).
#include "Optimizer.h"

#include "../../include/ba_b200.h"

#include <cstdio>
#include <cstring>
#include <unordered_map>

namespace {

void toRowMajor4(const Matrix4f &m, float out[16]) {
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) out[4 * r + c] = m(r, c);
}
void toRowMajor3(const Matrix3f &m, float out[9]) {
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) out[3 * r + c] = m(r, c);
}
Matrix4f fromRowMajor4(const float in[16]) {
    Matrix4f m;
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) m(r, c) = in[4 * r + c];
    return m;
}

// Optimizer.cpp:262-267: angle-axis -> R (double) -> float extrinsics -> inverse -> Frame::setPose
void writeBackPose(const std::shared_ptr<Frame> &frame, const double cam[6]) {
    float pose[16];
    ba_cam_to_pose(cam, pose);
    frame->setPose(fromRowMajor4(pose));
}

}  // namespace

BAOptimizer::BAOptimizer() : m_nIterations(3), m_nMaxItPerBA(200) {}

BAOptimizer::~BAOptimizer() {
    if (m_handle) ba_destroy(m_handle);
}

bool BAOptimizer::fail(int status) {
    m_lastStatus = status;
    if (status != BA_OK) {
        m_lastError = ba_last_error();
        std::fprintf(stderr, "[ba_b200] %s\n", m_lastError.c_str());   // the reference's signatures are void: report, do not throw
        return true;
    }
    return false;
}

ba_handle *BAOptimizer::handle() {
    if (!m_handle) {
        ba_options opt;
        ba_default_options(&opt);
        if (fail(ba_create(&opt, m_device, &m_handle))) m_handle = nullptr;
    }
    return m_handle;
}

// ------------------------------------------------------------------------------------------------------------
// pruneCorrespondences (Optimizer.cpp:6-79), batched over frames
// ------------------------------------------------------------------------------------------------------------
void BAOptimizer::pruneCorrespondences(const std::vector<std::shared_ptr<Frame>> &frames, bool considerOutlier) {
    const int F = (int)frames.size();
    if (F == 0) return;
    std::vector<float> pose(16 * (size_t)F), K(9 * (size_t)F);
    std::vector<int32_t> ptr(F + 1, 0);
    for (int f = 0; f < F; ++f) ptr[f + 1] = ptr[f] + frames[f]->getKeypointCount();
    const size_t n = (size_t)ptr[F];
    std::vector<float> xy(2 * n), xyz(3 * n, 0.f), mn(n, 0.f), mx(n, 0.f);
    std::vector<int32_t> octave(n, 0);
    std::vector<uint8_t> has(n, 0), out(n, 0);
    for (int f = 0; f < F; ++f) {
        const std::shared_ptr<Frame> &frame = frames[f];
        toRowMajor4(frame->getPose(), &pose[16 * (size_t)f]);
        toRowMajor3(frame->getIntrinsics(), &K[9 * (size_t)f]);
        const int cnt = ptr[f + 1] - ptr[f];
        for (int i = 0; i < cnt; ++i) {
            const size_t k = (size_t)ptr[f] + i;
            const cv::KeyPoint *kp = frame->getKeypoint(i);
            xy[2 * k] = kp->pt.x; xy[2 * k + 1] = kp->pt.y;
            octave[k] = kp->octave;
            out[k] = frame->isOutlier(i) ? 1 : 0;
            std::shared_ptr<MapPoint> mp = frame->getMapPoint(i);
            if (mp) {
                has[k] = 1;
                const Vector3f p = mp->getPosition();
                xyz[3 * k] = p(0); xyz[3 * k + 1] = p(1); xyz[3 * k + 2] = p(2);
                mn[k] = mp->getMinDistance(); mx[k] = mp->getMaxDistance();
            }
        }
    }
    if (n == 0) return;
    std::vector<uint8_t> before(out);
    if (fail(ba_prune_frames(m_device, F, pose.data(), K.data(), ptr.data(), xy.data(), octave.data(), xyz.data(), mn.data(),
                             mx.data(), has.data(), out.data(), considerOutlier ? 1 : 0)))
        return;
    for (int f = 0; f < F; ++f)
        for (int i = 0, cnt = ptr[f + 1] - ptr[f]; i < cnt; ++i) {
            const size_t k = (size_t)ptr[f] + i;
            if (out[k] == before[k]) continue;
            if (out[k]) frames[f]->setOutlier(i); else frames[f]->setInlier(i);
        }
}

void BAOptimizer::pruneCorrespondences(std::shared_ptr<Frame> frame, bool considerOutlier) {
    pruneCorrespondences(std::vector<std::shared_ptr<Frame>>{frame}, considerOutlier);
}

// ------------------------------------------------------------------------------------------------------------
// one ceres::Problem build + ceres::Solve + write-back (Global: Optimizer.cpp:229-271; Local: :557-599)
// ------------------------------------------------------------------------------------------------------------
bool BAOptimizer::solveOnce(const std::vector<std::shared_ptr<Frame>> &freeFrames, const std::vector<std::shared_ptr<Frame>> &fixedFrames,
                            const std::vector<std::shared_ptr<MapPoint>> &points, bool fixedOnlyKnownPoints) {
    (void)fixedOnlyKnownPoints;   // fixed frames only ever constrain points of the list (Optimizer.cpp:679)
    const int nFree = (int)freeFrames.size(), nFixed = (int)fixedFrames.size();
    const int Nc = nFree + nFixed, Np = (int)points.size();
    if (Nc == 0) return true;
    // Optimizer.cpp:282-289: float xyz -> double parameter block per map point
    std::unordered_map<const MapPoint *, int> pointIndex;
    pointIndex.reserve((size_t)Np * 2);
    std::vector<double> pt(3 * (size_t)std::max(Np, 1));
    for (int i = 0; i < Np; ++i) {
        const Vector3f p = points[i]->getPosition();
        pt[3 * (size_t)i] = (double)p(0); pt[3 * (size_t)i + 1] = (double)p(1); pt[3 * (size_t)i + 2] = (double)p(2);
        pointIndex.emplace(points[i].get(), i);
    }
    std::vector<double> cam(6 * (size_t)Nc);
    std::vector<uint8_t> camFixed(Nc, 0);
    std::vector<float> extr(16 * (size_t)Nc), K(9 * (size_t)Nc);
    std::vector<int32_t> obsCam, obsPt;
    std::vector<float> obsUv;
    for (int i = 0; i < Nc; ++i) {
        const bool isFixedList = i >= nFree;
        const std::shared_ptr<Frame> &frame = isFixedList ? fixedFrames[i - nFree] : freeFrames[i];
        float pose[16];
        toRowMajor4(frame->getPose(), pose);
        // Optimizer.cpp:296-299: extr = pose.inverse() (float), rotation -> angle-axis (double), translation -> double
        ba_pose_to_cam(pose, &cam[6 * (size_t)i], &extr[16 * (size_t)i]);
        toRowMajor3(frame->getIntrinsics(), &K[9 * (size_t)i]);
        // frame ID 0 is the world anchor: point-only residuals with its float extrinsics baked in (Optimizer.cpp:315-321)
        camFixed[i] = (isFixedList || frame->getID() == 0) ? 1 : 0;
        const std::vector<std::shared_ptr<MapPoint>> visible = frame->getMapPoints();
        for (int j = 0; j < (int)visible.size(); ++j) {
            const std::shared_ptr<MapPoint> &mp = visible[j];
            if (mp == nullptr || frame->isOutlier(j)) continue;
            auto it = pointIndex.find(mp.get());
            // Global BA precondition: every referenced point is in the map (the reference dereferences end() otherwise,
            // Optimizer.cpp:320/328); Local BA: fixed frames constrain only the listed points (Optimizer.cpp:679).
            if (it == pointIndex.end()) continue;
            const Vector2f uv = mp->getCorresponding2DKeyPointPosition(frame);
            obsCam.push_back(i);
            obsPt.push_back(it->second);
            obsUv.push_back(uv(0));
            obsUv.push_back(uv(1));
        }
    }
    const int No = (int)obsCam.size();
    ba_handle *h = handle();
    if (!h) return false;
    ba_options opt;
    ba_default_options(&opt);                       // configureSolver (Optimizer.cpp:80-90) + Ceres defaults
    opt.max_num_iterations = (int32_t)m_nMaxItPerBA;
    if (fail(ba_set_options(h, &opt))) return false;
    if (fail(ba_set_problem(h, Nc, cam.data(), camFixed.data(), extr.data(), K.data(), Np, pt.data(), No, obsCam.data(),
                            obsPt.data(), obsUv.data())))
        return false;
    ba_summary sm;
    if (fail(ba_solve(h, &sm))) return false;
    m_lastTermination = sm.termination_type;
    m_lastInitialCost = sm.initial_cost;
    m_lastFinalCost = sm.final_cost;
    m_lastNumIterations = sm.num_iterations;
    if (fail(ba_get_cameras(h, cam.data()))) return false;
    if (Np && fail(ba_get_points(h, pt.data()))) return false;
    // Optimizer.cpp:252-258: every listed map point gets its (double -> float) position back
    for (int i = 0; i < Np; ++i) {
        Vector3f p;
        p(0) = (float)pt[3 * (size_t)i]; p(1) = (float)pt[3 * (size_t)i + 1]; p(2) = (float)pt[3 * (size_t)i + 2];
        points[i]->setPosition(p);
    }
    // Optimizer.cpp:260-267: every FREE-LIST frame gets its pose back (frame 0 too: angle-axis round trip of its own pose)
    for (int i = 0; i < nFree; ++i) writeBackPose(freeFrames[i], &cam[6 * (size_t)i]);
    return true;
}

// ------------------------------------------------------------------------------------------------------------
// the whole outer loop of Global / Local BA (Optimizer.cpp:222-271, 550-603) in ONE device-resident call: the object graph is
// walked once (problem arrays as in solveOnce + one record per keypoint of every frame for pruneCorrespondences), then
// ba_set_problem + ba_solve_outer, then ONE write-back through the setters.  Between outer iterations nothing crosses PCIe and
// no index is rebuilt; a concurrent viewer thread sees the final state instead of every intermediate one.
// ------------------------------------------------------------------------------------------------------------
bool BAOptimizer::solveOuter(const std::vector<std::shared_ptr<Frame>> &freeFrames, const std::vector<std::shared_ptr<Frame>> &fixedFrames,
                             const std::vector<std::shared_ptr<MapPoint>> &points, unsigned nOuter) {
    const int nFree = (int)freeFrames.size(), nFixed = (int)fixedFrames.size();
    const int Nc = nFree + nFixed, Np = (int)points.size();
    if (Nc == 0 || nOuter == 0) return true;
    std::unordered_map<const MapPoint *, int> pointIndex;
    pointIndex.reserve((size_t)Np * 2);
    std::vector<double> pt(3 * (size_t)std::max(Np, 1));
    for (int i = 0; i < Np; ++i) {
        const Vector3f p = points[i]->getPosition();                       // Optimizer.cpp:282-289
        pt[3 * (size_t)i] = (double)p(0); pt[3 * (size_t)i + 1] = (double)p(1); pt[3 * (size_t)i + 2] = (double)p(2);
        pointIndex.emplace(points[i].get(), i);
    }
    std::vector<double> cam(6 * (size_t)Nc);
    std::vector<uint8_t> camFixed(Nc, 0), writeBack(Nc, 0);
    std::vector<float> pose(16 * (size_t)Nc), extr(16 * (size_t)Nc), K(9 * (size_t)Nc);
    std::vector<int32_t> obsCam, obsPt, framePtr(Nc + 1, 0), kpOctave, kpPt, kpObs;
    std::vector<float> obsUv, kpXy, kpXyz, kpMin, kpMax;
    std::vector<uint8_t> kpHas, kpOut;
    for (int i = 0; i < Nc; ++i) {
        const bool isFixedList = i >= nFree;
        const std::shared_ptr<Frame> &frame = isFixedList ? fixedFrames[i - nFree] : freeFrames[i];
        toRowMajor4(frame->getPose(), &pose[16 * (size_t)i]);
        ba_pose_to_cam(&pose[16 * (size_t)i], &cam[6 * (size_t)i], &extr[16 * (size_t)i]);       // Optimizer.cpp:296-299
        toRowMajor3(frame->getIntrinsics(), &K[9 * (size_t)i]);
        camFixed[i] = (isFixedList || frame->getID() == 0) ? 1 : 0;                              // Optimizer.cpp:315-321
        writeBack[i] = isFixedList ? 0 : 1;                                                      // Optimizer.cpp:260-267
        const std::vector<std::shared_ptr<MapPoint>> visible = frame->getMapPoints();
        const int cnt = frame->getKeypointCount();
        for (int j = 0; j < cnt; ++j) {
            const cv::KeyPoint *kp = frame->getKeypoint(j);
            const std::shared_ptr<MapPoint> mp = j < (int)visible.size() ? visible[j] : nullptr;
            const bool out = frame->isOutlier(j);
            kpXy.push_back(kp->pt.x); kpXy.push_back(kp->pt.y);
            kpOctave.push_back(kp->octave);
            kpOut.push_back(out ? 1 : 0);
            int pIdx = -1, oIdx = -1;
            float xyz[3] = {0.f, 0.f, 0.f}, mn = 0.f, mx = 0.f;
            if (mp) {
                const Vector3f p = mp->getPosition();
                xyz[0] = p(0); xyz[1] = p(1); xyz[2] = p(2);
                mn = mp->getMinDistance(); mx = mp->getMaxDistance();
                auto it = pointIndex.find(mp.get());
                if (it != pointIndex.end()) {
                    pIdx = it->second;
                    if (!out) {                                          // a residual block of the problem (Optimizer.cpp:308-328)
                        const Vector2f uv = mp->getCorresponding2DKeyPointPosition(frame);
                        oIdx = (int)obsCam.size();
                        obsCam.push_back(i); obsPt.push_back(pIdx);
                        obsUv.push_back(uv(0)); obsUv.push_back(uv(1));
                    }
                }
            }
            kpHas.push_back(mp ? 1 : 0);
            kpPt.push_back(pIdx); kpObs.push_back(oIdx);
            kpXyz.push_back(xyz[0]); kpXyz.push_back(xyz[1]); kpXyz.push_back(xyz[2]);
            kpMin.push_back(mn); kpMax.push_back(mx);
        }
        framePtr[i + 1] = (int32_t)kpOut.size();
    }
    const int No = (int)obsCam.size();
    ba_handle *h = handle();
    if (!h) return false;
    ba_options opt;
    ba_default_options(&opt);                       // configureSolver (Optimizer.cpp:80-90) + Ceres defaults
    opt.max_num_iterations = (int32_t)m_nMaxItPerBA;
    if (fail(ba_set_options(h, &opt))) return false;
    if (fail(ba_set_problem(h, Nc, cam.data(), camFixed.data(), extr.data(), K.data(), Np, pt.data(), No, obsCam.data(),
                            obsPt.data(), obsUv.data())))
        return false;
    std::vector<uint8_t> before(kpOut);
    std::vector<float> pt32(3 * (size_t)std::max(Np, 1));
    ba_summary sm;
    if (fail(ba_solve_outer(h, (int32_t)nOuter, writeBack.data(), pose.data(), framePtr.data(), kpXy.data(), kpOctave.data(), kpPt.data(),
                            kpXyz.data(), kpMin.data(), kpMax.data(), kpHas.data(), kpObs.data(), kpOut.data(), pt32.data(), &sm)))
        return false;
    m_lastTermination = sm.termination_type;
    m_lastInitialCost = sm.initial_cost;
    m_lastFinalCost = sm.final_cost;
    m_lastNumIterations = sm.num_iterations;
    for (int i = 0; i < Np; ++i) {                                       // Optimizer.cpp:252-258
        Vector3f p;
        p(0) = pt32[3 * (size_t)i]; p(1) = pt32[3 * (size_t)i + 1]; p(2) = pt32[3 * (size_t)i + 2];
        points[i]->setPosition(p);
    }
    for (int i = 0; i < nFree; ++i) freeFrames[i]->setPose(fromRowMajor4(&pose[16 * (size_t)i]));      // Optimizer.cpp:260-267
    for (int i = 0; i < Nc; ++i) {                                       // pruneCorrespondences' verdicts (Optimizer.cpp:270, 597-603)
        const std::shared_ptr<Frame> &frame = i >= nFree ? fixedFrames[i - nFree] : freeFrames[i];
        for (int j = 0, cnt = framePtr[i + 1] - framePtr[i]; j < cnt; ++j) {
            const size_t k = (size_t)framePtr[i] + j;
            if (kpOut[k] == before[k]) continue;
            if (kpOut[k]) frame->setOutlier(j); else frame->setInlier(j);
        }
    }
    return true;
}

// ------------------------------------------------------------------------------------------------------------
// GlobalBAOptimizerAngles (Optimizer.cpp:216-277)
// ------------------------------------------------------------------------------------------------------------
void GlobalBAOptimizerAngles::optimizeCamerasAndMapPoints(SceneMap *optimizedMap, bool eraseOutliers, int currentFrameID) {
    auto frames = optimizedMap->getKeyFrames();
    auto points = optimizedMap->getMapPoints();
    solveOuter(frames, {}, points, m_nIterations);   // m_nIterations x { Solve, write-back, prune } (Optimizer.cpp:222-271), on the device
    if (eraseOutliers) SfMHelper::getInstance()->eraseOutlier(optimizedMap, currentFrameID);
}

// ------------------------------------------------------------------------------------------------------------
// LocalBAOptimizerAngles (Optimizer.cpp:500-610)
// ------------------------------------------------------------------------------------------------------------
void LocalBAOptimizerAngles::optimizeCamerasAndMapPoints(SceneMap *optimizedMap, std::shared_ptr<Frame> currentFrame, bool eraseOutliers) {
    std::vector<std::shared_ptr<Frame>> localFrames = currentFrame->getBestCovisibilityFrames(10);
    localFrames.push_back(currentFrame);
    // distinct local map points in first-seen order (Optimizer.cpp:513-529; a hash set replaces the O(n^2) std::find)
    std::vector<std::shared_ptr<MapPoint>> localMapPoints;
    std::unordered_map<const MapPoint *, int> seenPoint;
    for (auto &frame : localFrames) {
        const auto mps = frame->getMapPoints();
        for (int j = 0; j < (int)mps.size(); ++j) {
            const std::shared_ptr<MapPoint> &mp = mps[j];
            if (mp && !frame->isOutlier(j) && !mp->isInvalid() && seenPoint.emplace(mp.get(), 1).second) localMapPoints.push_back(mp);
        }
    }
    // keyframes that observe a local point but are not optimised add constraints as fixed cameras (Optimizer.cpp:531-546)
    std::vector<std::shared_ptr<Frame>> fixedFrames;
    std::unordered_map<const Frame *, int> seenFrame;
    for (auto &f : localFrames) seenFrame.emplace(f.get(), 1);
    for (auto &mp : localMapPoints) {
        const auto observers = mp->getObservingKeyframes();
        for (auto &kv : observers) {
            if (kv.first->isOutlier((int)kv.second) || !kv.first->isKeyFrame()) continue;
            if (seenFrame.emplace(kv.first.get(), 1).second) fixedFrames.push_back(kv.first);
        }
    }
    solveOuter(localFrames, fixedFrames, localMapPoints, m_nIterations);   // Optimizer.cpp:550-603, on the device
    if (eraseOutliers) SfMHelper::getInstance()->eraseOutlier(optimizedMap, currentFrame->getID());
}

// ------------------------------------------------------------------------------------------------------------
// MotionOnlyBAOptimizerAngles (Optimizer.cpp:417-498)
// ------------------------------------------------------------------------------------------------------------
void MotionOnlyBAOptimizerAngles::optimizeCameraPose(std::shared_ptr<Frame> frame) {
    optimizeCameraPoses(std::vector<std::shared_ptr<Frame>>{frame});
}

void MotionOnlyBAOptimizerAngles::optimizeCameraPoses(const std::vector<std::shared_ptr<Frame>> &frames) {
    const int B = (int)frames.size();
    if (B == 0) return;
    ba_options opt;
    ba_default_options(&opt);
    opt.max_num_iterations = (int32_t)m_nMaxItPerBA;
    // The whole outer loop (Optimizer.cpp:422-456: build, Solve, write-back, pruneCorrespondences(frame, true)) runs on the
    // device: the object graph is walked ONCE into per-keypoint arrays, ba_solve_pose_only_outer iterates, and poses and outlier
    // bits are written back through the setters at the end.  A map point whose registered keypoint position differs from the
    // frame's own keypoint (two keypoints matched to one map point) cannot be expressed per keypoint: such batches take the
    // per-iteration host loop below, which calls the same kernels one outer iteration at a time.
    std::vector<int32_t> ptr(B + 1, 0);
    for (int b = 0; b < B; ++b) ptr[b + 1] = ptr[b] + frames[b]->getKeypointCount();
    const size_t n = (size_t)ptr[B];
    std::vector<float> pose(16 * (size_t)B), K(9 * (size_t)B), xy(2 * n), xyz(3 * n, 0.f), mn(n, 0.f), mx(n, 0.f);
    std::vector<int32_t> octave(n, 0);
    std::vector<uint8_t> has(n, 0), out(n, 0);
    bool perKeypoint = true;
    for (int b = 0; b < B && perKeypoint; ++b) {
        const std::shared_ptr<Frame> &frame = frames[b];
        toRowMajor4(frame->getPose(), &pose[16 * (size_t)b]);
        toRowMajor3(frame->getIntrinsics(), &K[9 * (size_t)b]);
        const auto points = frame->getMapPoints();
        for (int j = 0, cnt = ptr[b + 1] - ptr[b]; j < cnt; ++j) {
            const size_t k = (size_t)ptr[b] + j;
            const cv::KeyPoint *kp = frame->getKeypoint(j);
            xy[2 * k] = kp->pt.x; xy[2 * k + 1] = kp->pt.y;
            octave[k] = kp->octave;
            out[k] = frame->isOutlier(j) ? 1 : 0;
            const std::shared_ptr<MapPoint> mp = j < (int)points.size() ? points[j] : nullptr;
            if (!mp) continue;
            has[k] = 1;
            const Vector3f p = mp->getPosition();
            xyz[3 * k] = p(0); xyz[3 * k + 1] = p(1); xyz[3 * k + 2] = p(2);
            mn[k] = mp->getMinDistance(); mx[k] = mp->getMaxDistance();
            const Vector2f reg = mp->getCorresponding2DKeyPointPosition(frame);
            if (reg(0) != kp->pt.x || reg(1) != kp->pt.y) { perKeypoint = false; break; }
        }
    }
    if (perKeypoint) {
        std::vector<uint8_t> before(out);
        std::vector<ba_summary> sms(B);
        if (fail(ba_solve_pose_only_outer(&opt, m_device, B, pose.data(), K.data(), ptr.data(), xy.data(), octave.data(), xyz.data(),
                                          mn.data(), mx.data(), has.data(), out.data(), (int32_t)m_nIterations, sms.data(), nullptr, nullptr)))
            return;
        if (m_nIterations > 0) {
            m_lastTermination = sms[0].termination_type;
            m_lastInitialCost = sms[0].initial_cost;
            m_lastFinalCost = sms[0].final_cost;
            m_lastNumIterations = sms[0].num_iterations;
            for (int b = 0; b < B; ++b) frames[b]->setPose(fromRowMajor4(&pose[16 * (size_t)b]));   // Optimizer.cpp:448-453
        }
        for (int b = 0; b < B; ++b)
            for (int j = 0, cnt = ptr[b + 1] - ptr[b]; j < cnt; ++j) {
                const size_t k = (size_t)ptr[b] + j;
                if (out[k] == before[k]) continue;
                if (out[k]) frames[b]->setOutlier(j); else frames[b]->setInlier(j);
            }
        return;
    }
    for (unsigned it = 0; it < m_nIterations; ++it) {
        std::vector<double> cam(6 * (size_t)B);
        std::vector<float> K(9 * (size_t)B), Xh, uv;
        std::vector<int32_t> ptr(B + 1, 0);
        for (int b = 0; b < B; ++b) {
            const std::shared_ptr<Frame> &frame = frames[b];
            float pose[16], extr[16];
            toRowMajor4(frame->getPose(), pose);
            ba_pose_to_cam(pose, &cam[6 * (size_t)b], extr);                  // Optimizer.cpp:462-465
            toRowMajor3(frame->getIntrinsics(), &K[9 * (size_t)b]);
            const auto points = frame->getMapPoints();
            for (int j = 0; j < (int)points.size(); ++j) {
                const std::shared_ptr<MapPoint> &mp = points[j];
                if (mp == nullptr || frame->isOutlier(j)) continue;           // Optimizer.cpp:472
                const Vector3f p = mp->getPosition();                         // baked in as float homogeneous (:482)
                const Vector2f kp = mp->getCorresponding2DKeyPointPosition(frame);
                Xh.push_back(p(0)); Xh.push_back(p(1)); Xh.push_back(p(2)); Xh.push_back(1.0f);
                uv.push_back(kp(0)); uv.push_back(kp(1));
            }
            ptr[b + 1] = (int32_t)(uv.size() / 2);
        }
        std::vector<ba_summary> sms(B);
        if (fail(ba_solve_pose_only_batch(&opt, m_device, B, cam.data(), K.data(), ptr.data(), Xh.data(), uv.data(), nullptr,
                                          sms.data(), nullptr, 0)))
            return;
        m_lastTermination = sms[0].termination_type;
        m_lastInitialCost = sms[0].initial_cost;
        m_lastFinalCost = sms[0].final_cost;
        m_lastNumIterations = sms[0].num_iterations;
        for (int b = 0; b < B; ++b) writeBackPose(frames[b], &cam[6 * (size_t)b]);   // Optimizer.cpp:448-453
        pruneCorrespondences(frames, true);                                            // Optimizer.cpp:455
    }
}
