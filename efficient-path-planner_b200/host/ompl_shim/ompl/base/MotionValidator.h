// Minimal stand-in for <ompl/base/MotionValidator.h> (both checkMotion overloads pure virtual,
// valid_/invalid_ counters present as in OMPL; the reference never touches them).
#pragma once
#include <memory>
#include <utility>

#include "State.h"
namespace ompl {
namespace base {
class SpaceInformation;
using SpaceInformationPtr = std::shared_ptr<SpaceInformation>;
class MotionValidator {
public:
    explicit MotionValidator(SpaceInformation *si) : si_(si) {}
    explicit MotionValidator(const SpaceInformationPtr &si) : si_(si.get()) {}
    virtual ~MotionValidator() = default;
    virtual bool checkMotion(const State *s1, const State *s2) const = 0;
    virtual bool checkMotion(const State *s1, const State *s2, std::pair<State *, double> &lastValid) const = 0;
    unsigned int getValidMotionCount() const { return valid_; }
    unsigned int getInvalidMotionCount() const { return invalid_; }

protected:
    SpaceInformation *si_;
    mutable unsigned int valid_ = 0;
    mutable unsigned int invalid_ = 0;
};
using MotionValidatorPtr = std::shared_ptr<MotionValidator>;
} // namespace base
} // namespace ompl
