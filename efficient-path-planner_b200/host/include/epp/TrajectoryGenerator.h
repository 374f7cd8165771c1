// TrajectoryGenerator.h — poly_traj::generateTrajectory (external/poly_traj/include/poly_traj/trajectory_generator.h,
// src/trajectory_generator.cpp:12-100) on the GPU: Nfabian segment times (epp_segment_times), the batched linear
// min-snap solve (epp_minsnap) and the three evaluateRange passes (epp_sample_trajectories) behind the reference's
// signature.  result: N x 10 rows [px vx ax py vy ay pz vz az t + startTimeOffset].
#pragma once

#include <vector>

#include "LinAlg.h"

namespace poly_traj {

bool generateTrajectory(const std::vector<epp::Vector3d> &waypoints, double v_max, double a_max, double sampling_intervall,
                        const double startTimeOffset, const epp::Vector3d &initialVel, const epp::Vector3d &initialAcc,
                        epp::MatrixXd &result);

} // namespace poly_traj
