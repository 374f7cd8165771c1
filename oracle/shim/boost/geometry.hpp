// oracle/shim/boost/geometry.hpp — TEST INFRASTRUCTURE.  Stand-in for the part of Boost.Geometry the
// reference uses (include/Types.h:5-16, src/World.cpp:65,75,83,109,143): model::box, an rtree that is a
// plain list (query results are a set; the reference only asks "does any candidate collide"), and the two
// query predicates with Boost's published semantics for cartesian points and boxes:
//   index::contains(point)   -> value box CONTAINS the point  == within(point, box): strict, the boundary
//                               is excluded (min < p < max on every axis)
//   index::intersects(box)   -> closed boxes share at least one point (min_a <= max_b && min_b <= max_a)
#pragma once
#include <algorithm>
#include <cstddef>
#include <utility>
#include <vector>

#define BOOST_GEOMETRY_REGISTER_POINT_3D(Point, CoordinateType, CoordinateSystem, Field0, Field1, Field2)

namespace boost {
namespace geometry {
namespace cs {
struct cartesian {};
} // namespace cs
namespace model {
template <typename P>
class box {
public:
    box() = default;
    box(const P &mn, const P &mx) : min_(mn), max_(mx) {}
    const P &min_corner() const { return min_; }
    const P &max_corner() const { return max_; }
    P &min_corner() { return min_; }
    P &max_corner() { return max_; }

private:
    P min_, max_;
};
} // namespace model

namespace index {
template <std::size_t N>
struct quadratic {};

template <typename P>
struct contains_pred {
    P p;
};
template <typename B>
struct intersects_pred {
    B b;
};
template <typename P>
contains_pred<P> contains(const P &p)
{
    return contains_pred<P>{p};
}
template <typename B>
intersects_pred<B> intersects(const B &b)
{
    return intersects_pred<B>{b};
}

template <typename Value, typename Params>
class rtree {
public:
    void insert(const Value &v) { values_.push_back(v); }
    // removes one value equal to v (box corners and payload compare equal)
    std::size_t remove(const Value &v)
    {
        for (std::size_t i = 0; i < values_.size(); ++i)
            if (equal(values_[i], v)) {
                values_.erase(values_.begin() + i);
                return 1;
            }
        return 0;
    }
    void clear() { values_.clear(); }
    std::size_t size() const { return values_.size(); }

    template <typename P, typename Out>
    std::size_t query(const contains_pred<P> &q, Out out) const
    {
        std::size_t n = 0;
        for (const Value &v : values_) {
            const auto &mn = v.first.min_corner();
            const auto &mx = v.first.max_corner();
            bool in = true;
            for (int d = 0; d < 3; ++d) in = in && (mn(d) < q.p(d)) && (q.p(d) < mx(d));
            if (in) {
                *out++ = v;
                ++n;
            }
        }
        return n;
    }
    template <typename B, typename Out>
    std::size_t query(const intersects_pred<B> &q, Out out) const
    {
        std::size_t n = 0;
        for (const Value &v : values_) {
            const auto &mn = v.first.min_corner();
            const auto &mx = v.first.max_corner();
            bool hit = true;
            for (int d = 0; d < 3; ++d)
                hit = hit && (q.b.min_corner()(d) <= mx(d)) && (mn(d) <= q.b.max_corner()(d));
            if (hit) {
                *out++ = v;
                ++n;
            }
        }
        return n;
    }

private:
    static bool equal(const Value &a, const Value &b)
    {
        if (!(a.second == b.second)) return false;
        for (int d = 0; d < 3; ++d)
            if (a.first.min_corner()(d) != b.first.min_corner()(d) || a.first.max_corner()(d) != b.first.max_corner()(d))
                return false;
        return true;
    }
    std::vector<Value> values_;
};
} // namespace index
} // namespace geometry
} // namespace boost
