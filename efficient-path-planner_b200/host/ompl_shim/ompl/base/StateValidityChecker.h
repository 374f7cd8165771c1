// Minimal stand-in for <ompl/base/StateValidityChecker.h>.
#pragma once
#include <memory>

#include "State.h"
namespace ompl {
namespace base {
class SpaceInformation;
using SpaceInformationPtr = std::shared_ptr<SpaceInformation>;
class StateValidityChecker {
public:
    explicit StateValidityChecker(SpaceInformation *si) : si_(si) {}
    explicit StateValidityChecker(const SpaceInformationPtr &si) : si_(si.get()) {}
    virtual ~StateValidityChecker() = default;
    virtual bool isValid(const State *state) const = 0;

protected:
    SpaceInformation *si_;
};
using StateValidityCheckerPtr = std::shared_ptr<StateValidityChecker>;
} // namespace base
} // namespace ompl
