// LinAlg.h — the three dense types the reference's public signatures use (Eigen::Vector3d,
// Eigen::VectorXd, Eigen::MatrixXd; include/World.h:41-92, include/PathPlanner.h:30-80 of the
// reference).  With -DEPP_USE_EIGEN they ARE the Eigen types; without Eigen (this image has none)
// they are minimal stand-ins with the same element access, so call sites compile unchanged.
// No arithmetic lives here: every check runs on the GPU behind the C ABI.
#pragma once

#ifdef EPP_USE_EIGEN
#include <Eigen/Dense>
namespace epp {
using Vector3d = Eigen::Vector3d;
using VectorXd = Eigen::VectorXd;
using MatrixXd = Eigen::MatrixXd; // column-major
} // namespace epp
#else
#include <cstddef>
#include <initializer_list>
#include <stdexcept>
#include <vector>

namespace epp {

struct Vector3d {
    double v[3] = {0.0, 0.0, 0.0};
    Vector3d() = default;
    Vector3d(double x, double y, double z) : v{x, y, z} {}
    double &operator()(int i) { return v[i]; }
    double operator()(int i) const { return v[i]; }
    double &operator[](int i) { return v[i]; }
    double operator[](int i) const { return v[i]; }
    double x() const { return v[0]; }
    double y() const { return v[1]; }
    double z() const { return v[2]; }
    const double *data() const { return v; }
    double *data() { return v; }
    static constexpr int size() { return 3; }
};

class VectorXd {
public:
    VectorXd() = default;
    explicit VectorXd(std::size_t n) : d_(n, 0.0) {}
    VectorXd(std::initializer_list<double> l) : d_(l) {}
    VectorXd(const double *p, std::size_t n) : d_(p, p + n) {}
    double &operator()(std::size_t i) { return d_[i]; }
    double operator()(std::size_t i) const { return d_[i]; }
    double &operator[](std::size_t i) { return d_[i]; }
    double operator[](std::size_t i) const { return d_[i]; }
    std::size_t size() const { return d_.size(); }
    const double *data() const { return d_.data(); }
    double *data() { return d_.data(); }

private:
    std::vector<double> d_;
};

// Column-major like Eigen::MatrixXd: element (r, c) at data()[c * rows() + r].
class MatrixXd {
public:
    MatrixXd() = default;
    MatrixXd(std::size_t rows, std::size_t cols) : r_(rows), c_(cols), d_(rows * cols, 0.0) {}
    // from a row-major buffer (what numpy hands over)
    static MatrixXd fromRowMajor(const double *p, std::size_t rows, std::size_t cols)
    {
        MatrixXd m(rows, cols);
        for (std::size_t r = 0; r < rows; ++r)
            for (std::size_t c = 0; c < cols; ++c) m(r, c) = p[r * cols + c];
        return m;
    }
    double &operator()(std::size_t r, std::size_t c) { return d_[c * r_ + r]; }
    double operator()(std::size_t r, std::size_t c) const { return d_[c * r_ + r]; }
    std::size_t rows() const { return r_; }
    std::size_t cols() const { return c_; }
    const double *data() const { return d_.data(); }
    double *data() { return d_.data(); }
    VectorXd row(std::size_t r) const
    {
        VectorXd v(c_);
        for (std::size_t c = 0; c < c_; ++c) v(c) = (*this)(r, c);
        return v;
    }

private:
    std::size_t r_ = 0, c_ = 0;
    std::vector<double> d_;
};

} // namespace epp
#endif
