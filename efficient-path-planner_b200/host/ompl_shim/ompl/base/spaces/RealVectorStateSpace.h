// Minimal stand-in for <ompl/base/spaces/RealVectorStateSpace.h>: StateType::values is a double*
// to `dimension` doubles, exactly the layout the reference validators read
// (src/StateValidator.cpp:9-11, src/MotionValidator.cpp:10-14).
#pragma once
#include <vector>

#include "../StateSpace.h"
namespace ompl {
namespace base {
class RealVectorBounds {
public:
    explicit RealVectorBounds(unsigned int dim) : low(dim, 0.0), high(dim, 0.0) {}
    void setLow(unsigned int i, double v) { low[i] = v; }
    void setHigh(unsigned int i, double v) { high[i] = v; }
    void setLow(double v) { low.assign(low.size(), v); }
    void setHigh(double v) { high.assign(high.size(), v); }
    std::vector<double> low, high;
};

class RealVectorStateSpace : public StateSpace {
public:
    class StateType : public State {
    public:
        double operator[](unsigned int i) const { return values[i]; }
        double &operator[](unsigned int i) { return values[i]; }
        double *values = nullptr;
    };
    explicit RealVectorStateSpace(unsigned int dim = 0) : dimension_(dim), bounds_(dim) {}
    unsigned int getDimension() const override { return dimension_; }
    void setBounds(const RealVectorBounds &b) { bounds_ = b; }
    const RealVectorBounds &getBounds() const { return bounds_; }
    State *allocState() const override
    {
        auto *s = new StateType();
        s->values = new double[dimension_]();
        return s;
    }
    void freeState(State *state) const override
    {
        auto *s = static_cast<StateType *>(state);
        delete[] s->values;
        delete s;
    }

private:
    unsigned int dimension_;
    RealVectorBounds bounds_;
};
} // namespace base
} // namespace ompl
