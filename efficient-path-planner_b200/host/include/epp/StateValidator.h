// StateValidator.h — the reference's ompl::base::StateValidityChecker subclass
// (include/StateValidator.h:15-36, src/StateValidator.cpp:7-13): same class name, constructor and
// isValid override, so it installs with si->setStateValidityChecker(...) unchanged
// (src/PathPlanner.cpp:49).  Added: isValidBatch for sample batches (FMT's sampling phase, RRT*'s
// candidate sets), one kernel launch for the whole batch.
#pragma once

#include <cstdint>
#include <memory>
#include <vector>

#include <ompl/base/SpaceInformation.h>
#include <ompl/base/StateValidityChecker.h>
#include <ompl/base/spaces/RealVectorStateSpace.h>

#include "World.h"

class StateValidator : public ompl::base::StateValidityChecker {
public:
    StateValidator(const ompl::base::SpaceInformationPtr &si, const std::shared_ptr<World> &world, const bool canPassGate)
        : ompl::base::StateValidityChecker(si), world(world), canPassGate(canPassGate){};

    // World::checkPointValidity(values[0..2], canPassGate) — one state, on the GPU.
    virtual bool isValid(const ompl::base::State *state) const override;

    // Batched: maskWords needs World::maskWords(n) words; bit = valid.
    void isValidBatch(const ompl::base::State *const *states, std::size_t n, uint32_t *maskWords) const;
    // Same on a packed n x 3 row-major array.
    void isValidBatch(const double *xyz, std::size_t n, uint32_t *maskWords) const;
    std::vector<bool> isValidBatch(const std::vector<const ompl::base::State *> &states) const;

private:
    std::shared_ptr<World> world;
    bool canPassGate;
};
