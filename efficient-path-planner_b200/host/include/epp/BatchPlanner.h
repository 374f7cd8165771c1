// BatchPlanner.h — batch-native sampling planner that stands where the reference calls OMPL's RRT* /
// FMT (src/PathPlanner.cpp:106-139).  It asks the validators for whole batches instead of one state /
// one edge per virtual call:
//   1. N uniform samples in the world bounds + N/8 around start and goal -> StateValidator::isValidBatch (1 launch)
//   2. radius-neighbour edges among the valid samples   -> MotionValidator::checkMotionBatch (1 launch)
//   3. Dijkstra over the valid edges (host), doubling N until a path exists or the time limit ends
//   4. vertex reduction: all (i, j) shortcuts of the path -> checkMotionBatch (1 launch), then the
//      shortest chain of valid shortcuts (the batched analogue of PathSimplifier::reduceVertices, :138-139)
// The roadmap construction is FMT-like (fixed sample set, connection radius shrinking with N); it is a
// different planner from OMPL's and makes no claim of returning OMPL's paths.
#pragma once

#include <cstdint>
#include <vector>

#include "LinAlg.h"
#include "MotionValidator.h"
#include "StateValidator.h"

namespace epp {

struct BatchPlannerStats {
    int rounds = 0;
    std::size_t samples = 0;
    std::size_t validSamples = 0;
    std::size_t edgesChecked = 0;
    std::size_t validEdges = 0;
    std::size_t shortcutChecks = 0;
    double pathLength = 0.0;
    double seconds = 0.0;
};

class BatchPlanner {
public:
    BatchPlanner(const StateValidator &sv, const MotionValidator &mv, const Vector3d &lower, const Vector3d &upper)
        : sv_(sv), mv_(mv), lo_(lower), hi_(upper) {}

    void setInitialSamples(std::size_t n) { initialSamples_ = n; }
    void setMaxSamples(std::size_t n) { maxSamples_ = n; }
    void setSeed(uint64_t s) { seed_ = s; }
    // connection radius = max(range, gamma * (log n / n)^(1/3)) like FMT*/PRM*
    void setRange(double r) { range_ = r; }

    bool solve(const Vector3d &start, const Vector3d &goal, double timeLimit, std::vector<Vector3d> &path,
               BatchPlannerStats *stats = nullptr) const;

    // Batched vertex reduction of an existing path (every pairwise shortcut in one edge batch).
    std::vector<Vector3d> reduceVertices(const std::vector<Vector3d> &path, BatchPlannerStats *stats = nullptr) const;

private:
    const StateValidator &sv_;
    const MotionValidator &mv_;
    Vector3d lo_, hi_;
    std::size_t initialSamples_ = 2048;
    std::size_t maxSamples_ = 1u << 18;
    uint64_t seed_ = 0x5EED0001ull;
    double range_ = 0.5;
};

} // namespace epp
