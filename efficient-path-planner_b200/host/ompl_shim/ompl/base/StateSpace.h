// Minimal stand-in for <ompl/base/StateSpace.h>.
#pragma once
#include <memory>

#include "State.h"
namespace ompl {
namespace base {
class StateSpace {
public:
    virtual ~StateSpace() = default;
    template <class T>
    T *as()
    {
        return static_cast<T *>(this);
    }
    template <class T>
    const T *as() const
    {
        return static_cast<const T *>(this);
    }
    virtual unsigned int getDimension() const = 0;
    virtual State *allocState() const = 0;
    virtual void freeState(State *state) const = 0;
};
using StateSpacePtr = std::shared_ptr<StateSpace>;
} // namespace base
} // namespace ompl
