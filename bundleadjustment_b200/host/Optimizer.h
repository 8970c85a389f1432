// Optimizer.h -- drop-in replacements for the reference's BA optimizer classes
// (ba_project/src/ba/Optimizer.h:199-223 BAOptimizer, :242-255 GlobalBAOptimizerAngles, :268-277
// MotionOnlyBAOptimizerAngles, :279-289 LocalBAOptimizerAngles).  Same class names, same public signatures, same
// defaults (3 outer iterations, 200 LM iterations per ceres::Solve); underneath, every ceres::Problem build +
// ceres::Solve is one ba_set_problem + ba_solve on the GPU (include/ba_b200.h) and pruneCorrespondences is one
// batched kernel launch (ba_prune_frames).  No Ceres, Eigen or OpenCV needed.
//
// Data model: with -DBA_B200_REFERENCE_MODEL the reference's own Frame / MapPoint / SceneMap / SfMHelper headers are
// used (a maintainer's build); otherwise the accessor-compatible mocks in ba_model_mock.h.
#pragma once

#ifdef BA_B200_REFERENCE_MODEL
#include "../model/SceneMap.h"
#include "SfMHelper.h"
#else
#include "ba_model_mock.h"
#endif

#include <memory>
#include <string>
#include <vector>

struct ba_handle;   // include/ba_b200.h

/** BA optimizer - base class (Optimizer.h:199-223). */
class BAOptimizer {
public:
    BAOptimizer();
    ~BAOptimizer();
    BAOptimizer(const BAOptimizer &) = delete;
    BAOptimizer &operator=(const BAOptimizer &) = delete;

    void setNbOfIterations(unsigned nIterations) { m_nIterations = nIterations; }
    void setNbOfMaxItPerBA(unsigned nMaxItPerBA) { m_nMaxItPerBA = nMaxItPerBA; }

    // ---- additions (the reference ignores ceres::Solver::Summary, Optimizer.cpp:241-242; we keep it reachable) ----
    void setDevice(int device) { m_device = device; }
    int lastStatus() const { return m_lastStatus; }                 // 0 = BA_OK
    const std::string &lastError() const { return m_lastError; }
    int lastTerminationType() const { return m_lastTermination; }  // BA_CONVERGENCE / BA_NO_CONVERGENCE / BA_FAILURE
    double lastInitialCost() const { return m_lastInitialCost; }
    double lastFinalCost() const { return m_lastFinalCost; }
    int lastNumIterations() const { return m_lastNumIterations; }

protected:
    unsigned m_nIterations;
    unsigned m_nMaxItPerBA;
    int m_device = 0;
    int m_lastStatus = 0, m_lastTermination = 0, m_lastNumIterations = 0;
    double m_lastInitialCost = 0.0, m_lastFinalCost = 0.0;
    std::string m_lastError;
    ba_handle *m_handle = nullptr;      // lazily created, one per optimizer object

    /** pruneCorrespondences (Optimizer.cpp:6-79) for a batch of frames: one kernel launch, float32 semantics. */
    void pruneCorrespondences(const std::vector<std::shared_ptr<Frame>> &frames, bool considerOutlier);
    void pruneCorrespondences(std::shared_ptr<Frame> frame, bool considerOutlier = true);
    bool fail(int status);
    ba_handle *handle();

    /** Shared by Global and Local BA: the WHOLE outer loop (nOuter x { solve, write-back, prune }) in one device-resident call
     *  (ba_set_problem + ba_solve_outer); the object graph is read once and written once. */
    bool solveOuter(const std::vector<std::shared_ptr<Frame>> &freeFrames, const std::vector<std::shared_ptr<Frame>> &fixedFrames,
                    const std::vector<std::shared_ptr<MapPoint>> &points, unsigned nOuter);
    /** One outer iteration on the host side of the boundary: flatten (free frames | fixed frames | points) -> solve -> write back
     *  (kept for callers that interleave their own work between outer iterations). */
    bool solveOnce(const std::vector<std::shared_ptr<Frame>> &freeFrames, const std::vector<std::shared_ptr<Frame>> &fixedFrames,
                   const std::vector<std::shared_ptr<MapPoint>> &points, bool fixedOnlyKnownPoints);
};

/** Global BA over all keyframes and all map points, angle-axis poses (Optimizer.cpp:216-333). */
class GlobalBAOptimizerAngles : public BAOptimizer {
public:
    GlobalBAOptimizerAngles() {}
    void optimizeCamerasAndMapPoints(SceneMap *optimizedMap, bool eraseOutliers, int currentFrameID);
};

/** Frame-to-frame "BA" pose estimation: one 6-dof block, landmarks fixed, Huber always on (Optimizer.cpp:417-498). */
class MotionOnlyBAOptimizerAngles : public BAOptimizer {
public:
    MotionOnlyBAOptimizerAngles() {}
    void optimizeCameraPose(std::shared_ptr<Frame> frame);
    /** Addition: the same optimisation for many frames at once -- every outer iteration is ONE batched
     *  solve launch + ONE batched prune launch; results per frame are identical to calling optimizeCameraPose on each. */
    void optimizeCameraPoses(const std::vector<std::shared_ptr<Frame>> &frames);
};

/** Windowed BA: current frame + 10 best covisible frames free, other observers fixed (Optimizer.cpp:500-698). */
class LocalBAOptimizerAngles : public BAOptimizer {
public:
    LocalBAOptimizerAngles() {}
    void optimizeCamerasAndMapPoints(SceneMap *optimizedMap, std::shared_ptr<Frame> currentFrame, bool eraseOutliers);
};
