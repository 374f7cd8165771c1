// Minimal stand-in for <ompl/base/SpaceInformation.h>: holds the space and the two validators and
// forwards isValid / checkMotion to them, which is how OMPL planners reach the reference's validators
// (src/PathPlanner.cpp:48-51).
#pragma once
#include <memory>
#include <stdexcept>

#include "MotionValidator.h"
#include "StateSpace.h"
#include "StateValidityChecker.h"
namespace ompl {
namespace base {
class SpaceInformation {
public:
    explicit SpaceInformation(StateSpacePtr space) : space_(std::move(space)) {}
    const StateSpacePtr &getStateSpace() const { return space_; }
    void setStateValidityChecker(const StateValidityCheckerPtr &svc) { svc_ = svc; }
    void setMotionValidator(const MotionValidatorPtr &mv) { mv_ = mv; }
    const StateValidityCheckerPtr &getStateValidityChecker() const { return svc_; }
    const MotionValidatorPtr &getMotionValidator() const { return mv_; }
    void setup()
    {
        if (!svc_ || !mv_) throw std::runtime_error("SpaceInformation::setup: validators not set");
        setup_ = true;
    }
    bool isSetup() const { return setup_; }
    bool isValid(const State *s) const { return svc_->isValid(s); }
    bool checkMotion(const State *a, const State *b) const { return mv_->checkMotion(a, b); }
    State *allocState() const { return space_->allocState(); }
    void freeState(State *s) const { space_->freeState(s); }

private:
    StateSpacePtr space_;
    StateValidityCheckerPtr svc_;
    MotionValidatorPtr mv_;
    bool setup_ = false;
};
using SpaceInformationPtr = std::shared_ptr<SpaceInformation>;
} // namespace base
} // namespace ompl
