// oracle/shim_poly/ref_poly_capi.cpp — TEST INFRASTRUCTURE.  C entry points over the reference's OWN
// mav_trajectory_generation::Polynomial / Segment / Trajectory / Vertex, compiled unmodified from
// /root/reference/external/poly_traj/src into oracle/_ref/libepp_refpoly.so together with this file
// (oracle/Makefile target `refpoly`).  Pins the oracle's restatement of the sampler rows a13/a14
// (SURVEY.md §8a): Horner evaluation, Segment::evaluate, Trajectory::evaluateRange and the
// estimateSegmentTimesNfabian heuristic.  rpoly_ak1.cpp (Jenkins-Traub roots, not on the sampling path) is
// not compiled; its two entry points are defined here and abort if ever reached.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "mav_trajectory_generation/polynomial.h"
#include "mav_trajectory_generation/polynomial_optimization_linear.h"
#include "poly_traj/trajectory_generator.h"
#include "mav_trajectory_generation/rpoly/rpoly_ak1.h"
#include "mav_trajectory_generation/segment.h"
#include "mav_trajectory_generation/trajectory.h"
#include "mav_trajectory_generation/vertex.h"

namespace mtg = mav_trajectory_generation;

namespace mav_trajectory_generation {
int findLastNonZeroCoeff(const Eigen::VectorXd &)
{
    std::fprintf(stderr, "ref_poly_capi: rpoly is not part of this build\n");
    std::abort();
}
bool findRootsJenkinsTraub(const Eigen::VectorXd &, Eigen::VectorXcd *)
{
    std::fprintf(stderr, "ref_poly_capi: rpoly is not part of this build\n");
    std::abort();
}
} // namespace mav_trajectory_generation

namespace {
mtg::Trajectory make_trajectory(const double *coeffs, const double *durations, int K, int D, int N)
{
    mtg::Segment::Vector segments;
    for (int k = 0; k < K; ++k) {
        mtg::Segment seg(N, D);
        for (int d = 0; d < D; ++d) {
            Eigen::VectorXd c(N);
            for (int j = 0; j < N; ++j) c[j] = coeffs[((size_t)k * D + d) * N + j];
            seg[d] = mtg::Polynomial(N, c);
        }
        seg.setTime(durations[k]);
        segments.push_back(seg);
    }
    mtg::Trajectory traj;
    traj.setSegments(segments); // accumulates max_time_ as in trajectory.h:54-71
    return traj;
}
} // namespace

extern "C" {

// Polynomial::evaluate(t, derivative), polynomial.h:136-150
double refpoly_horner(const double *coeffs, int N, double t, int derivative)
{
    Eigen::VectorXd c(N);
    for (int j = 0; j < N; ++j) c[j] = coeffs[j];
    return mtg::Polynomial(N, c).evaluate(t, derivative);
}

double refpoly_base_coefficient(int derivative, int j) { return mtg::Polynomial::base_coefficients_(derivative, j); }

// Trajectory::getMaxTime() of the trajectory built from (coeffs [K][D][N], durations [K])
double refpoly_max_time(const double *coeffs, const double *durations, int K, int D, int N)
{
    return make_trajectory(coeffs, durations, K, D, N).getMaxTime();
}

// The call of poly_traj::generateTrajectory (trajectory_generator.cpp:60-69):
// evaluateRange(getMinTime(), getMaxTime(), dt, derivative, &samples, &times).  Writes up to max_samples rows
// of D values and their sampling times; returns the number of samples the reference produced.
long refpoly_evaluate_range(const double *coeffs, const double *durations, int K, int D, int N, double dt,
                            int derivative, double *out, double *out_times, long max_samples)
{
    const mtg::Trajectory traj = make_trajectory(coeffs, durations, K, D, N);
    std::vector<Eigen::VectorXd> samples;
    std::vector<double> times;
    traj.evaluateRange(traj.getMinTime(), traj.getMaxTime(), dt, derivative, &samples, &times);
    const long n = (long)samples.size();
    for (long i = 0; i < n && i < max_samples; ++i) {
        for (int d = 0; d < D; ++d) out[(size_t)i * D + d] = samples[(size_t)i](d);
        if (out_times) out_times[i] = times[(size_t)i];
    }
    return n;
}

// Trajectory::evaluate(t, derivative), trajectory.cpp:48-79 (single time)
void refpoly_evaluate(const double *coeffs, const double *durations, int K, int D, int N, double t, int derivative,
                      double *out)
{
    const Eigen::VectorXd v = make_trajectory(coeffs, durations, K, D, N).evaluate(t, derivative);
    for (int d = 0; d < D; ++d) out[d] = v(d);
}

// estimateSegmentTimesNfabian over position-only vertices (vertex.cpp:272-289), as generateTrajectory's
// estimateSegmentTimes(vertices, v_max, a_max) does with its default constant 6.5.
void refpoly_segment_times(const double *waypoints, int n, int D, double v_max, double a_max, double *out)
{
    mtg::Vertex::Vector vertices;
    for (int i = 0; i < n; ++i) {
        mtg::Vertex v(D);
        Eigen::VectorXd p(D);
        for (int d = 0; d < D; ++d) p[d] = waypoints[(size_t)i * D + d];
        v.addConstraint(mtg::derivative_order::POSITION, p);
        vertices.push_back(v);
    }
    const std::vector<double> t = mtg::estimateSegmentTimes(vertices, v_max, a_max);
    for (size_t i = 0; i < t.size(); ++i) out[i] = t[i];
}

} // extern "C"

// ---- the linear min-snap solve ------------------------------------------------------------
// The reference's OWN formulation code (PolynomialOptimization<10>::setupFromVertices / solveLinear, compiled
// unmodified from polynomial_optimization_linear_impl.h) over the stand-in linear algebra of shim_poly/Eigen.
extern "C" {

// poly_traj::generateTrajectory exactly as OnlineTrajGenerator calls it (trajectory_generator.cpp:12-100).
// rows: row-major n x 10, capacity `cap` rows.  Returns the row count or -1.
long refpoly_generate_trajectory(const double *waypoints, int n_wp, double v_max, double a_max, double dt, double offset,
                                 const double *vel3, const double *acc3, double *rows, long cap)
{
    std::vector<Eigen::Vector3d> wp;
    for (int i = 0; i < n_wp; ++i) wp.push_back(Eigen::Vector3d(waypoints[3 * i], waypoints[3 * i + 1], waypoints[3 * i + 2]));
    Eigen::MatrixXd result;
    if (!poly_traj::generateTrajectory(wp, v_max, a_max, dt, offset, Eigen::Vector3d(vel3[0], vel3[1], vel3[2]),
                                       Eigen::Vector3d(acc3[0], acc3[1], acc3[2]), result))
        return -1;
    if (result.rows() > cap) return -1;
    for (long r = 0; r < (long)result.rows(); ++r)
        for (int c = 0; c < 10; ++c) rows[10 * r + c] = result(r, c);
    return (long)result.rows();
}

// Coefficients for GIVEN segment times: the vertex pattern of generateTrajectory (start: position, velocity,
// acceleration given, jerk = snap = 0; interior: position; end: position, derivatives 0) handed to the reference's
// optimiser.  coeffs: [K][3][10].
int refpoly_minsnap(const double *waypoints, int n_wp, const double *times, const double *vel3, const double *acc3,
                    double *coeffs)
{
    const int D = 3, snap = mtg::derivative_order::SNAP;
    mtg::Vertex::Vector vertices;
    mtg::Vertex start(D);
    Eigen::VectorXd sc(3 * D);
    for (int d = 0; d < D; ++d) {
        sc[d] = waypoints[d];
        sc[D + d] = vel3[d];
        sc[2 * D + d] = acc3[d];
    }
    start.makeStartOrEnd(sc, snap);
    vertices.push_back(start);
    for (int i = 1; i + 1 < n_wp; ++i) {
        mtg::Vertex v(D);
        v.addConstraint(mtg::derivative_order::POSITION, Eigen::Vector3d(waypoints[3 * i], waypoints[3 * i + 1], waypoints[3 * i + 2]));
        vertices.push_back(v);
    }
    mtg::Vertex end(D);
    end.makeStartOrEnd(Eigen::Vector3d(waypoints[3 * (n_wp - 1)], waypoints[3 * (n_wp - 1) + 1], waypoints[3 * (n_wp - 1) + 2]), snap);
    vertices.push_back(end);
    std::vector<double> t(times, times + (n_wp - 1));
    mtg::PolynomialOptimization<10> opt(D);
    if (!opt.setupFromVertices(vertices, t, snap)) return -1;
    if (!opt.solveLinear()) return -1;
    mtg::Segment::Vector segments;
    opt.getSegments(&segments);
    for (int k = 0; k < n_wp - 1; ++k)
        for (int d = 0; d < D; ++d) {
            const Eigen::VectorXd c = segments[k][d].getCoefficients(0);
            for (int j = 0; j < 10; ++j) coeffs[(k * D + d) * 10 + j] = c[j];
        }
    return 0;
}

} // extern "C"
