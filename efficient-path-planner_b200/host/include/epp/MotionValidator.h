// MotionValidator.h — the reference's ompl::base::MotionValidator subclass
// (include/MotionValidator.h:13-32, src/MotionValidator.cpp:8-23): same class name, constructor and
// both checkMotion overrides.  The two-state form is the reference's exact segment-vs-OBB slab test
// (World::checkRayValid) — the reference does not discretise motions.  The three-argument form keeps
// the reference's behaviour: it prints "Not implemented error" to stderr and returns true.
// Added: checkMotionBatch for edge batches.
#pragma once

#include <cstdint>
#include <memory>
#include <utility>
#include <vector>

#include <ompl/base/MotionValidator.h>
#include <ompl/base/SpaceInformation.h>
#include <ompl/base/spaces/RealVectorStateSpace.h>

#include "World.h"

class MotionValidator : public ompl::base::MotionValidator {
public:
    MotionValidator(const ompl::base::SpaceInformationPtr &si, const std::shared_ptr<World> &world, const bool canPassGate)
        : ompl::base::MotionValidator(si), world(world), canPassGate(canPassGate){};

    virtual bool checkMotion(const ompl::base::State *s1, const ompl::base::State *s2) const override;
    virtual bool checkMotion(const ompl::base::State *s1, const ompl::base::State *s2,
                             std::pair<ompl::base::State *, double> &lastValid) const override;

    void checkMotionBatch(const ompl::base::State *const *s1, const ompl::base::State *const *s2, std::size_t n,
                          uint32_t *maskWords) const;
    void checkMotionBatch(const double *starts, const double *ends, std::size_t n, uint32_t *maskWords) const;

private:
    std::shared_ptr<World> world;
    const bool canPassGate;
};
