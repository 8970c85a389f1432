// ba_model_mock.h -- stand-ins for the reference's data model (ba_project/src/model/{Frame,MapPoint,SceneMap}.h,
// utils/Eigen.h, cv::KeyPoint) exposing the SAME accessors the BA optimisation path touches.
//
// Why mocks: the real headers pull in Eigen and OpenCV, neither of which exists in this image.  The drop-in
// optimizer classes (Optimizer.h / Optimizer.cpp next to this file) only use element access `m(r, c)` / `v(i)` and
// the getters / setters below, all of which the real types provide with identical spelling, so compiling with
// -DBA_B200_REFERENCE_MODEL against the reference's own headers needs no source change.
//
// Semantics mirrored from the reference (file:line of ba_project/src/model/):
//   Frame::getPose/setPose        camera->world 4x4 float, mutex-guarded            Frame.cpp:18-23, 87-91
//   Frame::getMapPoints           index-aligned with keypoints, nullptr = none     Frame.cpp:105-109
//   Frame::isOutlier/setOutlier   per-keypoint bit                                  Frame.cpp:179-183
//   Frame::getWorldPos            translation column of the pose                    Frame.cpp:191-195
//   MapPoint::getCorresponding2DKeyPointPosition  keypoint of the observing index   MapPoint.cpp:29-37
//   SceneMap::getKeyFrames        ordered by frame ID                               SceneMap.h:16-24, SceneMap.cpp:33-37
//   SceneMap::getMapPoints        ordered by shared_ptr ADDRESS                     SceneMap.h:25, SceneMap.cpp:39-43
#pragma once

#include <algorithm>
#include <map>
#include <memory>
#include <mutex>
#include <set>
#include <vector>

// ---- utils/Eigen.h stand-ins: fixed-size float matrices with Eigen's element-access spelling ----------------
namespace ba_mock {
template <int R, int C>
struct Mat {
    float v[R * C];     // row-major here; the adaptor never relies on the storage order
    Mat() { for (float &x : v) x = 0.f; }
    float &operator()(int r, int c) { return v[r * C + c]; }
    float operator()(int r, int c) const { return v[r * C + c]; }
    float &operator()(int i) { return v[i]; }
    float operator()(int i) const { return v[i]; }
    static Mat Identity() { Mat m; for (int i = 0; i < (R < C ? R : C); ++i) m(i, i) = 1.f; return m; }
};
}  // namespace ba_mock
typedef ba_mock::Mat<4, 4> Matrix4f;
typedef ba_mock::Mat<3, 3> Matrix3f;
typedef ba_mock::Mat<3, 1> Vector3f;
typedef ba_mock::Mat<2, 1> Vector2f;

namespace cv {
struct Point2f { float x = 0.f, y = 0.f; };
struct KeyPoint { Point2f pt; int octave = 0; };
}  // namespace cv

class Frame;

class MapPoint : public std::enable_shared_from_this<MapPoint> {
public:
    MapPoint(Vector3f worldPosition, std::shared_ptr<Frame> observingFrame, size_t indexInFrame)
        : m_worldPosition(worldPosition), m_referenceKeyframe(observingFrame) {
        if (observingFrame) m_observingKeyframes[observingFrame] = indexInFrame;     // MapPoint.cpp:11 guards on null
    }
    inline Vector2f getCorresponding2DKeyPointPosition(std::shared_ptr<Frame> frame);
    Vector3f getPosition() { std::lock_guard<std::mutex> l(m_pointMutex); return m_worldPosition; }
    void setPosition(Vector3f p) { std::lock_guard<std::mutex> l(m_pointMutex); m_worldPosition = p; }
    void addObservation(std::shared_ptr<Frame> f, size_t index) { std::lock_guard<std::mutex> l(m_pointMutex); m_observingKeyframes[f] = index; }
    void removeObservation(std::shared_ptr<Frame> f) {                               // MapPoint.cpp:79-104
        std::lock_guard<std::mutex> l(m_pointMutex);
        if (m_observingKeyframes.find(f) == m_observingKeyframes.end()) return;
        m_observingKeyframes.erase(f);
        if (f == m_referenceKeyframe) {
            if (!m_observingKeyframes.empty()) m_referenceKeyframe = m_observingKeyframes.begin()->first;
            else { m_referenceKeyframe = nullptr; m_invalid = true; }
        }
    }
    std::map<std::shared_ptr<Frame>, size_t> getObservingKeyframes() { std::lock_guard<std::mutex> l(m_pointMutex); return m_observingKeyframes; }
    int getNumObserved() { std::lock_guard<std::mutex> l(m_pointMutex); return (int)m_observingKeyframes.size(); }
    std::shared_ptr<Frame> getReferenceFrame() { return m_referenceKeyframe; }
    float getMinDistance() { return m_minDist; }
    float getMaxDistance() { return m_maxDist; }
    void setDistanceBounds(float mn, float mx) { m_minDist = mn; m_maxDist = mx; }   // (the reference derives them from the scale pyramid)
    inline void erase();
    bool isInvalid() { return m_invalid; }
    void computeDescriptor() {}          // appearance bookkeeping: not part of the BA path
    void computeViewingDirection() {}

private:
    Vector3f m_worldPosition;
    std::map<std::shared_ptr<Frame>, size_t> m_observingKeyframes;
    std::shared_ptr<Frame> m_referenceKeyframe;
    std::mutex m_pointMutex;
    float m_maxDist = 1e30f, m_minDist = 0.f;
    bool m_invalid = false;
};

class Frame : public std::enable_shared_from_this<Frame> {
public:
    Frame(Matrix4f pose, Matrix3f intrinsics, int frameID, double timeStamp = 0.0)
        : m_frameID(frameID), m_intrinsics(intrinsics), m_pose(pose), m_timeStamp(timeStamp) {}
    double getTimeStamp() { return m_timeStamp; }                                   // Frame.cpp:197-201
    void setPose(Matrix4f pose) { std::lock_guard<std::mutex> l(m_frameMutex); m_pose = pose; m_validPose = true; }
    Matrix4f getPose() { std::lock_guard<std::mutex> l(m_frameMutex); return m_pose; }
    Matrix3f getIntrinsics() { std::lock_guard<std::mutex> l(m_frameMutex); return m_intrinsics; }
    std::vector<std::shared_ptr<MapPoint>> getMapPoints() { std::lock_guard<std::mutex> l(m_frameMutex); return m_mapPoints; }
    std::shared_ptr<MapPoint> getMapPoint(int i) { std::lock_guard<std::mutex> l(m_frameMutex); return m_mapPoints[i]; }
    cv::KeyPoint *getKeypoint(int i) { std::lock_guard<std::mutex> l(m_frameMutex); return &m_keypoints[i]; }
    int getKeypointCount() { std::lock_guard<std::mutex> l(m_frameMutex); return (int)m_keypoints.size(); }
    void setKeypoints(std::vector<cv::KeyPoint> kps) {
        std::lock_guard<std::mutex> l(m_frameMutex);
        m_keypoints = kps;
        m_mapPoints.assign(kps.size(), nullptr);
        m_outlier.assign(kps.size(), false);
    }
    void addAssociatedMapPoint(int i, std::shared_ptr<MapPoint> p) { std::lock_guard<std::mutex> l(m_frameMutex); m_mapPoints[i] = p; }
    void eraseAssociatedMapPoint(int i) { std::lock_guard<std::mutex> l(m_frameMutex); m_mapPoints[i] = nullptr; }
    void setOutlier(int i) { std::lock_guard<std::mutex> l(m_frameMutex); m_outlier[i] = true; }
    void setInlier(int i) { std::lock_guard<std::mutex> l(m_frameMutex); m_outlier[i] = false; }
    bool isOutlier(int i) { std::lock_guard<std::mutex> l(m_frameMutex); return m_outlier[i]; }
    int getID() { std::lock_guard<std::mutex> l(m_frameMutex); return m_frameID; }
    Vector3f getWorldPos() {
        std::lock_guard<std::mutex> l(m_frameMutex);
        Vector3f p; p(0) = m_pose(0, 3); p(1) = m_pose(1, 3); p(2) = m_pose(2, 3);
        return p;
    }
    void setKeyFrame() { m_isKeyFrame = true; }
    bool isKeyFrame() { return m_isKeyFrame; }
    void setOrderedCovisibilityFrames(std::vector<std::shared_ptr<Frame>> f) { m_orderedCovisibilityFrames = f; }
    std::vector<std::shared_ptr<Frame>> getBestCovisibilityFrames(int n) {       // Frame.cpp:376-386
        std::lock_guard<std::mutex> l(m_frameMutex);
        if ((int)m_orderedCovisibilityFrames.size() <= n) return m_orderedCovisibilityFrames;
        return std::vector<std::shared_ptr<Frame>>(m_orderedCovisibilityFrames.begin(), m_orderedCovisibilityFrames.begin() + n);
    }

private:
    int m_frameID;
    Matrix3f m_intrinsics;
    std::vector<cv::KeyPoint> m_keypoints;
    std::vector<std::shared_ptr<MapPoint>> m_mapPoints;
    std::vector<bool> m_outlier;
    Matrix4f m_pose;
    double m_timeStamp = 0.0;
    bool m_validPose = false;
    std::vector<std::shared_ptr<Frame>> m_orderedCovisibilityFrames;
    std::mutex m_frameMutex;
    bool m_isKeyFrame = false;
};

inline Vector2f MapPoint::getCorresponding2DKeyPointPosition(std::shared_ptr<Frame> frame) {
    std::lock_guard<std::mutex> l(m_pointMutex);
    const size_t idx = m_observingKeyframes.find(frame)->second;     // precondition as in the reference (MapPoint.cpp:33)
    const cv::Point2f pt = frame->getKeypoint((int)idx)->pt;
    Vector2f v; v(0) = pt.x; v(1) = pt.y;
    return v;
}

inline void MapPoint::erase() {
    std::map<std::shared_ptr<Frame>, size_t> obs;
    { std::lock_guard<std::mutex> l(m_pointMutex); obs.swap(m_observingKeyframes); m_invalid = true; }
    for (auto &kv : obs) kv.first->eraseAssociatedMapPoint((int)kv.second);
}

class SceneMap {
protected:
    struct cmp { bool operator()(std::shared_ptr<Frame> a, std::shared_ptr<Frame> b) const { return a->getID() < b->getID(); } };
    std::set<std::shared_ptr<Frame>, cmp> m_keyframes;
    std::set<std::shared_ptr<MapPoint>> m_mapPoints;

public:
    void addKeyFrame(std::shared_ptr<Frame> f) { std::lock_guard<std::mutex> l(m_mapMutex); f->setKeyFrame(); m_keyframes.insert(f); }
    void addMapPoint(std::shared_ptr<MapPoint> p) { std::lock_guard<std::mutex> l(m_mapMutex); m_mapPoints.insert(p); }
    void eraseMapPoint(std::shared_ptr<MapPoint> p) { std::lock_guard<std::mutex> l(m_mapMutex); m_mapPoints.erase(p); }
    std::vector<std::shared_ptr<Frame>> getKeyFrames() { std::lock_guard<std::mutex> l(m_mapMutex); return {m_keyframes.begin(), m_keyframes.end()}; }
    std::vector<std::shared_ptr<MapPoint>> getMapPoints() { std::lock_guard<std::mutex> l(m_mapMutex); return {m_mapPoints.begin(), m_mapPoints.end()}; }
    int getMapPointCount() { std::lock_guard<std::mutex> l(m_mapMutex); return (int)m_mapPoints.size(); }
    int getKeyFrameCount() { std::lock_guard<std::mutex> l(m_mapMutex); return (int)m_keyframes.size(); }

private:
    std::mutex m_mapMutex;
};

// SfMHelper::eraseOutlier (ba_project/src/ba/SfMHelper.cpp:1079-1123), the optional post-step of the BA entry points.
class SfMHelper {
public:
    static SfMHelper *getInstance() { static SfMHelper s; return &s; }
    void eraseOutlier(SceneMap *map, int currentFrameID) {
        auto frames = map->getKeyFrames();
        auto points = map->getMapPoints();
        for (auto &f : frames) {
            const int n = f->getKeypointCount();
            for (int j = 0; j < n; ++j) {
                if (!f->isOutlier(j)) continue;
                std::shared_ptr<MapPoint> p = f->getMapPoint(j);
                if (p) { f->eraseAssociatedMapPoint(j); p->removeObservation(f); }
            }
        }
        for (auto &p : points) {
            if (p->isInvalid()) { map->eraseMapPoint(p); continue; }
            if (p->getNumObserved() <= 2 && (currentFrameID - p->getReferenceFrame()->getID()) >= 4) { p->erase(); map->eraseMapPoint(p); }
            else { p->computeDescriptor(); p->computeViewingDirection(); }
        }
    }
};
