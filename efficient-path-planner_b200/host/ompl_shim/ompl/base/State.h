// Minimal stand-in for <ompl/base/State.h> (OMPL 1.6 is not installed in this image).  Only what the
// validators and their tests touch; with real OMPL on the include path this directory is not used.
#pragma once
namespace ompl {
namespace base {
class State {
public:
    template <class T>
    const T *as() const
    {
        return static_cast<const T *>(this);
    }
    template <class T>
    T *as()
    {
        return static_cast<T *>(this);
    }
    State() = default;
    virtual ~State() = default;
    State(const State &) = delete;
    State &operator=(const State &) = delete;
};
} // namespace base
} // namespace ompl
