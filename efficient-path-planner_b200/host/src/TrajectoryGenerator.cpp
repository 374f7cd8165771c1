// TrajectoryGenerator.cpp — see TrajectoryGenerator.h.
#include "epp/TrajectoryGenerator.h"

#include <stdexcept>
#include <string>

#include "../../../include/epp_c_api.h"

namespace poly_traj {

namespace {
void check(int rc, const char *what)
{
    if (rc != EPP_OK) throw std::runtime_error(std::string(what) + ": " + epp_last_error());
}
} // namespace

bool generateTrajectory(const std::vector<epp::Vector3d> &waypoints, double v_max, double a_max, double sampling_intervall,
                        const double startTimeOffset, const epp::Vector3d &initialVel, const epp::Vector3d &initialAcc,
                        epp::MatrixXd &result)
{
    if (waypoints.size() < 2) throw std::invalid_argument("At least two waypoints are required"); // :21-25
    const int n = (int)waypoints.size(), K = n - 1, N = 10;
    std::vector<double> wp(3 * (size_t)n), T((size_t)K), coeffs((size_t)K * 3 * N);
    for (int i = 0; i < n; ++i)
        for (int d = 0; d < 3; ++d) wp[3 * (size_t)i + d] = waypoints[(size_t)i](d);
    check(epp_segment_times(wp.data(), n, v_max, a_max, T.data()), "epp_segment_times"); // :45
    const double va[6] = {initialVel(0), initialVel(1), initialVel(2), initialAcc(0), initialAcc(1), initialAcc(2)};
    check(epp_minsnap(wp.data(), T.data(), va, 1, K, coeffs.data()), "epp_minsnap"); // :46-48
    double total = 0.0;
    for (double t : T) total += t;
    const size_t ld = ((size_t)(total / sampling_intervall) + 8 + 31) / 32 * 32;
    std::vector<double> out(ld * 10);
    uint32_t rows = 0;
    check(epp_sample_trajectories(coeffs.data(), T.data(), &sampling_intervall, &startTimeOffset, 1, K, N, ld, out.data(),
                                  &rows),
          "epp_sample_trajectories"); // :56-97
    if (rows > ld) throw std::runtime_error("generateTrajectory: more rows than estimated");
    result = epp::MatrixXd(rows, 10);
    for (int c = 0; c < 10; ++c)
        for (uint32_t r = 0; r < rows; ++r) result(r, c) = out[(size_t)c * ld + r];
    return true;
}

} // namespace poly_traj
