// OnlineTrajGenerator.cpp — see OnlineTrajGenerator.h.  Line numbers cite src/OnlineTrajGenerator.cpp of the reference.
#include "epp/OnlineTrajGenerator.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <iostream>
#include <limits>
#include <stdexcept>

#include "epp/TrajectoryGenerator.h"

namespace {

epp::Vector3d axpy(const epp::Vector3d &c, double s, const epp::Vector3d &n)
{
    return epp::Vector3d(c(0) + s * n(0), c(1) + s * n(1), c(2) + s * n(2));
}

epp::Vector3d position(const epp::MatrixXd &traj, std::size_t row) { return epp::Vector3d(traj(row, 0), traj(row, 3), traj(row, 6)); }

// rows [r0, r1) of a trajectory matrix
epp::MatrixXd rowRange(const epp::MatrixXd &m, std::size_t r0, std::size_t r1)
{
    const std::size_t n = r1 > r0 ? r1 - r0 : 0;
    epp::MatrixXd out(n, m.cols());
    for (std::size_t c = 0; c < (std::size_t)m.cols(); ++c)
        for (std::size_t r = 0; r < n; ++r) out(r, c) = m(r0 + r, c);
    return out;
}

} // namespace

OnlineTrajGenerator::OnlineTrajGenerator(const epp::Vector3d start, const epp::Vector3d goal,
                                         const epp::MatrixXd &nominalGatePositionAndType,
                                         const epp::MatrixXd &nominalObstaclePosition, const std::string &configPath)
    : configParser(std::make_shared<ConfigParser>(configPath)), nominalGatePositionAndType(nominalGatePositionAndType),
      nominalObstaclePosition(nominalObstaclePosition),
      pathPlanner(nominalGatePositionAndType, nominalObstaclePosition, configParser),
      // include/OnlineTrajGenerator.h:132 writes to "path_segments" in the working directory; EPP_PATH_SEGMENTS_DIR moves it
      pathWriter(std::getenv("EPP_PATH_SEGMENTS_DIR") ? std::getenv("EPP_PATH_SEGMENTS_DIR") : "path_segments")
{
    for (std::size_t i = 0; i < (std::size_t)nominalGatePositionAndType.rows(); ++i) // :20-30
        pathWriter.updateGatePos((int)i, nominalGatePositionAndType.row(i));
    for (std::size_t i = 0; i < (std::size_t)nominalObstaclePosition.rows(); ++i)
        pathWriter.updateObstaclePos((int)i, nominalObstaclePosition.row(i));
    const double offset = configParser->getPathPlannerProperties().checkpointGateOffset; // :32-46
    checkpoints.push_back(start);
    for (std::size_t i = 0; i < (std::size_t)nominalGatePositionAndType.rows(); ++i) {
        epp::Vector3d center, normal;
        getGateCenterAndNormal(nominalGatePositionAndType.row(i), center, normal);
        checkpoints.push_back(axpy(center, -offset, normal));
        checkpoints.push_back(axpy(center, offset, normal));
    }
    checkpoints.push_back(goal);
    pathWriter.writeCheckpoints(checkpoints);
}

void OnlineTrajGenerator::getGateCenterAndNormal(const epp::VectorXd &gate, epp::Vector3d &center, epp::Vector3d &normal) const
{
    const double height = configParser->getObjectPropertiesByTypeId((int)gate(6)).height; // :49-69
    center = epp::Vector3d(gate(0), gate(1), gate(2) + height);
    const double nx = -std::sin(gate(5)), ny = std::cos(gate(5));
    const double len = std::sqrt(nx * nx + ny * ny);
    normal = epp::Vector3d(nx / len, ny / len, 0.0);
}

bool OnlineTrajGenerator::checkGatePassed(const epp::Vector3d &pos1, const epp::Vector3d &pos2, const int gateId) const
{
    const epp::VectorXd gate = nominalGatePositionAndType.row((std::size_t)gateId); // :229-258
    epp::Vector3d center, normal;
    getGateCenterAndNormal(gate, center, normal);
    const double c = std::cos(gate(5)), s = std::sin(gate(5));
    auto toGate = [&](const epp::Vector3d &p) {
        const double tx = p(0) - center(0), ty = p(1) - center(1), tz = p(2) - center(2);
        return epp::Vector3d(c * tx - s * ty, s * tx + c * ty, tz);
    };
    const epp::Vector3d g1 = toGate(pos1), g2 = toGate(pos2);
    if (g1(1) < 0 && g2(1) > 0) {
        const double mx = (g1(0) + g2(0)) / 2, mz = (g1(2) + g2(2)) / 2;
        return std::fabs(mx) <= 0.425 && std::fabs(mz) <= 0.425;
    }
    return false;
}

epp::MatrixXd OnlineTrajGenerator::generate(const std::vector<epp::Vector3d> &waypoints, double offset,
                                            const epp::Vector3d &vel, const epp::Vector3d &acc) const
{
    const TrajectoryGeneratorProperties &tg = configParser->getTrajectoryGeneratorProperties();
    if (tg.type != "snap") { // :94-118: "spline" and "optimal" are not built
        std::cerr << "Trajectory type not supported" << std::endl;
        throw std::runtime_error("Trajectory type not supported");
    }
    epp::MatrixXd traj;
    poly_traj::generateTrajectory(waypoints, tg.maxVelocity, tg.maxAcceleration, tg.samplingInterval, offset, vel, acc, traj);
    return traj;
}

void OnlineTrajGenerator::preComputeTraj(const double takeoffTime)
{
    const double timeLimit = configParser->getPathPlannerProperties().timeLimitOffline; // :72-121
    pathSegments.clear();
    for (std::size_t i = 0; i + 1 < checkpoints.size(); i += 2) {
        std::vector<epp::Vector3d> path;
        if (!pathPlanner.planPath(checkpoints[i], checkpoints[i + 1], timeLimit, path)) throw std::runtime_error("Path not found");
        pathSegments.push_back(path);
    }
    const std::vector<epp::Vector3d> prunedPath = pathPlanner.includeGates2(pathSegments);
    pathWriter.writePath(prunedPath);
    plannedTraj = generate(prunedPath, takeoffTime, epp::Vector3d(0, 0, 0), epp::Vector3d(0, 0, 0));
}

bool OnlineTrajGenerator::updateGatePos(const int gateId, const epp::VectorXd &newPose, const epp::Vector3d &dronePos,
                                        const bool nextGateWithinRange, const double flightTime)
{
    if (!nextGateWithinRange) return false;                                        // :127-130
    if (gatesObservedWithinRange.count(gateId)) return false;                      // :132-135
    if (!pathPlanner.worldPtr->checkPointValidity(dronePos, false)) return false;  // :137-140
    gatesObservedWithinRange.insert(gateId);
    for (int k = 0; k < 6; ++k) nominalGatePositionAndType((std::size_t)gateId, k) = newPose(k); // :144
    pathPlanner.parseGatesAndObstacles(nominalGatePositionAndType, nominalObstaclePosition);
    pathWriter.updateGatePos(gateId, nominalGatePositionAndType.row((std::size_t)gateId));

    const std::size_t rows = plannedTraj.rows(), last = plannedTraj.cols() - 1; // :149-165
    std::size_t startIdx = 0, endIdx = 0;
    double best = std::numeric_limits<double>::infinity();
    for (std::size_t r = 0; r < rows; ++r) {
        const double d = std::fabs(plannedTraj(r, last) - flightTime);
        if (d < best) best = d, startIdx = r;
    }
    const std::size_t cpNext = std::min<std::size_t>(2 * (std::size_t)gateId + 3, checkpoints.size() - 1);
    best = std::numeric_limits<double>::infinity();
    for (std::size_t r = 0; r < rows; ++r) {
        const epp::Vector3d p = position(plannedTraj, r);
        const double dx = p(0) - checkpoints[cpNext](0), dy = p(1) - checkpoints[cpNext](1), dz = p(2) - checkpoints[cpNext](2);
        const double d = std::sqrt(dx * dx + dy * dy + dz * dz);
        if (d < best) best = d, endIdx = r;
    }
    const epp::MatrixXd lookahead = rowRange(plannedTraj, startIdx, std::max(endIdx, startIdx));

    bool passing = false, valid = false; // :167-199
    for (std::size_t i = 0; i + 1 < (std::size_t)lookahead.rows() && !passing; ++i)
        passing = checkGatePassed(position(lookahead, i), position(lookahead, i + 1), gateId);
    if (!passing) {
        std::cout << "Current trajectory not passing next gate. Must be recomputed. No check for collision requried" << std::endl;
    } else {
        std::cout << "Current trajectory is passing next gate" << std::endl;
        valid = pathPlanner.checkTrajectoryValidity(lookahead, configParser->getPathPlannerProperties().minDistCheckTrajCollision);
        std::cout << (valid ? "Checked trajectory. It is not colliding, no need to recompute"
                            : "Checked trajectory. It is colliding, recomputing trajectory")
                  << std::endl;
    }
    if (valid && passing) return false;
    if (trajectoryCurrentlyUpdating) { // :206-210
        std::cerr << "Call to update trajectory, while previous update is still going on";
        throw std::runtime_error("Call to update trajectory, while previous update is still going on");
    }
    trajectoryCurrentlyUpdating = true;
    try {
        recomputeTraj(gateId, dronePos, flightTime);
    } catch (...) {
        trajectoryCurrentlyUpdating = false;
        throw;
    }
    trajectoryCurrentlyUpdating = false;
    return true;
}

void OnlineTrajGenerator::recomputeTraj(const int gateId, const epp::Vector3d &, const double flightTime)
{
    const PathPlannerProperties &pp = configParser->getPathPlannerProperties(); // :260-421
    const std::size_t pre = 2 * (std::size_t)gateId + 1, post = pre + 1, next = pre + 2;
    epp::Vector3d center, normal;
    getGateCenterAndNormal(nominalGatePositionAndType.row((std::size_t)gateId), center, normal);
    checkpoints[pre] = axpy(center, -pp.checkpointGateOffset, normal);
    checkpoints[post] = axpy(center, pp.checkpointGateOffset, normal);
    pathWriter.writeCheckpoints(checkpoints);
    double advanced = flightTime;
    if (pp.advanceForCalculation) advanced += pp.timeLimitOnline + 0.01;
    const std::size_t rows = plannedTraj.rows(), last = plannedTraj.cols() - 1;
    std::size_t idx = rows;
    for (std::size_t r = 0; r < rows; ++r)
        if (plannedTraj(r, last) > advanced) {
            idx = r;
            break;
        }
    if (idx + 1 >= rows) return;
    const epp::Vector3d p = position(plannedTraj, idx + 1);
    const epp::Vector3d v(plannedTraj(idx + 1, 1), plannedTraj(idx + 1, 4), plannedTraj(idx + 1, 7));
    const epp::Vector3d a(plannedTraj(idx + 1, 2), plannedTraj(idx + 1, 5), plannedTraj(idx + 1, 8));
    if (!pathPlanner.worldPtr->checkPointValidity(p, pp.canPassGate)) {
        std::cout << "Advanced trajectory does not end at valid position. No recomputation and hope for best" << std::endl;
        return;
    }
    std::vector<epp::Vector3d> prePath, postPath;
    if (!pathPlanner.planPath(p, checkpoints[pre], pp.timeLimitOnline, prePath)) throw std::runtime_error("Pre path not found. Exiting");
    if (!pathPlanner.planPath(checkpoints[post], checkpoints[next], pp.timeLimitOnline, postPath))
        throw std::runtime_error("Post segment path not found. Exiting");
    pathSegments[(std::size_t)gateId] = prePath;
    pathSegments[(std::size_t)gateId + 1] = postPath;
    const std::vector<std::vector<epp::Vector3d>> rest(pathSegments.begin() + gateId, pathSegments.end());
    const std::vector<epp::Vector3d> filled = pathPlanner.includeGates2(rest);
    pathWriter.writePath(filled);
    const epp::MatrixXd partPost = generate(filled, advanced, v, a);
    epp::MatrixXd joined(idx + partPost.rows(), plannedTraj.cols());
    for (std::size_t c = 0; c < (std::size_t)plannedTraj.cols(); ++c) {
        for (std::size_t r = 0; r < idx; ++r) joined(r, c) = plannedTraj(r, c);
        for (std::size_t r = 0; r < (std::size_t)partPost.rows(); ++r) joined(idx + r, c) = partPost(r, c);
    }
    plannedTraj = joined;
}

epp::VectorXd OnlineTrajGenerator::sampleTraj(const double currentTime) const
{
    if (plannedTraj.rows() == 0) throw std::runtime_error("No trajectory data available."); // :423-439
    const std::size_t last = plannedTraj.cols() - 1;
    std::size_t bestRow = 0;
    double best = std::numeric_limits<double>::infinity();
    for (std::size_t r = 0; r < (std::size_t)plannedTraj.rows(); ++r) {
        const double d = std::fabs(plannedTraj(r, last) - currentTime);
        if (d < best) best = d, bestRow = r;
    }
    return plannedTraj.row(bestRow);
}

double OnlineTrajGenerator::getTrajEndTime() const
{
    if (plannedTraj.rows() == 0) throw std::runtime_error("No trajectory data available."); // :441-450
    return plannedTraj(plannedTraj.rows() - 1, plannedTraj.cols() - 1);
}

epp::MatrixXd OnlineTrajGenerator::getPlannedTraj() const
{
    if (plannedTraj.rows() == 0) throw std::runtime_error("No trajectory data available."); // :452-460
    return plannedTraj;
}
