// PathPlanner.cpp — see PathPlanner.h.  Follows src/PathPlanner.cpp of the reference where cited.
#include "epp/PathPlanner.h"

#include <cmath>
#include <iostream>
#include <stdexcept>

#include "epp/BatchPlanner.h"

namespace ob = ompl::base;

PathPlanner::PathPlanner(const epp::MatrixXd &nominalGatePositionAndType, const epp::MatrixXd &nominalObstaclePosition,
                         std::shared_ptr<ConfigParser> configParser)
    : configParser(configParser)
{
    worldPtr = std::make_shared<World>(configParser);
    parseGatesAndObstacles(nominalGatePositionAndType, nominalObstaclePosition);

    // :37-44
    space = ob::StateSpacePtr(new ob::RealVectorStateSpace(3));
    ob::RealVectorBounds bounds(3);
    for (int i = 0; i < 3; i++) {
        bounds.setLow(i, configParser->getWorldProperties().lowerBound(i));
        bounds.setHigh(i, configParser->getWorldProperties().upperBound(i));
    }
    space->as<ob::RealVectorStateSpace>()->setBounds(bounds);

    // :46-51
    const bool canPassGates = configParser->getPathPlannerProperties().canPassGate;
    si = ob::SpaceInformationPtr(new ob::SpaceInformation(space));
    si->setStateValidityChecker(std::make_shared<StateValidator>(si, worldPtr, canPassGates));
    si->setMotionValidator(std::make_shared<MotionValidator>(si, worldPtr, canPassGates));
    si->setup();
}

void PathPlanner::parseGatesAndObstacles(const epp::MatrixXd &nominalGatePositionAndType,
                                         const epp::MatrixXd &nominalObstaclePosition)
{
    worldPtr->resetWorld();
    for (std::size_t i = 0; i < (std::size_t)nominalGatePositionAndType.rows(); i++) {
        epp::VectorXd gate = nominalGatePositionAndType.row(i);
        gate(2) = 0.0; // put all gates on the ground (:68)
        worldPtr->addGate(static_cast<int>(i), gate);
    }
    for (std::size_t i = 0; i < (std::size_t)nominalObstaclePosition.rows(); i++) {
        epp::VectorXd obstacle = nominalObstaclePosition.row(i);
        worldPtr->addObstacle(static_cast<int>(i), obstacle);
    }
}

bool PathPlanner::planPath(const epp::Vector3d &start, const epp::Vector3d &goal, const double timeLimit,
                           std::vector<epp::Vector3d> &resultPath) const
{
    if (!resultPath.empty()) { // :83-87
        resultPath.clear();
        std::cerr << "Result path not empty, clearing it" << std::endl;
    }
    const PathPlannerProperties &pp = configParser->getPathPlannerProperties();
    if (pp.planner != "rrt" && pp.planner != "fmt") { // :119-123
        std::cerr << "Unknown planner" << std::endl;
        throw std::runtime_error("Unknown planner");
    }
    const auto *sv = static_cast<const StateValidator *>(si->getStateValidityChecker().get());
    const auto *mv = static_cast<const MotionValidator *>(si->getMotionValidator().get());
    epp::BatchPlanner planner(*sv, *mv, configParser->getWorldProperties().lowerBound,
                              configParser->getWorldProperties().upperBound);
    planner.setRange(pp.range);
    if (pp.planner == "fmt") planner.setInitialSamples(static_cast<std::size_t>(std::max(pp.samplesFMT, 64)));
    return planner.solve(start, goal, timeLimit, resultPath);
}

void PathPlanner::updateGatePos(const int gateId, const epp::VectorXd &newPose)
{
    worldPtr->updateGatePosition(gateId, newPose);
}

bool PathPlanner::checkTrajectoryValidity(const epp::MatrixXd &trajectory, const double minDistance) const
{
    const std::size_t rows = trajectory.rows(), cols = trajectory.cols();
    if (rows == 0) return true; // the reference's loop body never runs
    if (cols < 7) throw std::runtime_error("trajectory matrix needs at least 7 columns (position in 0, 3, 6)");
    uint32_t valid = 0;
    // column-major storage: leading dimension = rows
    worldPtr->checkTrajectoriesBatch(trajectory.data(), 1, rows, rows, rows * cols, static_cast<int>(cols), minDistance,
                                     &valid, nullptr, nullptr);
    return (valid & 1u) != 0;
}

std::vector<bool> PathPlanner::checkTrajectoryValidityBatch(const std::vector<epp::MatrixXd> &trajectories,
                                                            const double minDistance) const
{
    std::vector<bool> out(trajectories.size(), true);
    // group equal-shaped matrices into one launch each (typical batches are homogeneous)
    std::size_t i = 0;
    while (i < trajectories.size()) {
        const std::size_t rows = trajectories[i].rows(), cols = trajectories[i].cols();
        std::size_t j = i;
        while (j < trajectories.size() && (std::size_t)trajectories[j].rows() == rows &&
               (std::size_t)trajectories[j].cols() == cols)
            ++j;
        if (rows > 0) {
            if (cols < 7) throw std::runtime_error("trajectory matrix needs at least 7 columns (position in 0, 3, 6)");
            const std::size_t n = j - i, span = 7 * rows; // only columns 0..6 are read
            std::vector<double> packed(n * span);
            for (std::size_t t = 0; t < n; ++t)
                std::copy(trajectories[i + t].data(), trajectories[i + t].data() + span, packed.begin() + t * span);
            std::vector<uint32_t> words(World::maskWords(n));
            worldPtr->checkTrajectoriesBatch(packed.data(), n, rows, rows, span, 7, minDistance, words.data(), nullptr,
                                             nullptr);
            for (std::size_t t = 0; t < n; ++t) out[i + t] = (words[t >> 5] >> (t & 31)) & 1u;
        }
        i = j;
    }
    return out;
}

std::vector<epp::Vector3d> PathPlanner::includeGates2(std::vector<std::vector<epp::Vector3d>> waypoints) const
{
    if (waypoints.empty()) return {};
    auto norm = [](const epp::Vector3d &a, const epp::Vector3d &b) {
        const double dx = a(0) - b(0), dy = a(1) - b(1), dz = a(2) - b(2);
        return std::sqrt(dx * dx + dy * dy + dz * dz);
    };
    // step 1: gate centres = midpoints between consecutive segments (:178-184)
    std::vector<epp::Vector3d> gateCenters;
    for (std::size_t s = 0; s + 1 < waypoints.size(); s++) {
        const epp::Vector3d &e = waypoints[s].back();
        const epp::Vector3d &b = waypoints[s + 1].front();
        gateCenters.push_back(epp::Vector3d((e(0) + b(0)) / 2, (e(1) + b(1)) / 2, (e(2) + b(2)) / 2));
    }
    // step 2 (:187-194)
    for (std::size_t i = 0; i < gateCenters.size(); i++) {
        waypoints[i].push_back(gateCenters[i]);
        waypoints[i + 1].insert(waypoints[i + 1].begin(), gateCenters[i]);
    }
    std::vector<epp::Vector3d> flat;
    const std::string method = configParser->getPathPlannerProperties().pathSimplification;
    for (const auto &segment : waypoints) {
        std::vector<epp::Vector3d> pruned;
        if (method == "none")
            pruned = segment;
        else if (method == "custom")
            pruned = pruneWaypoints(segment);
        else if (method == "ompl")
            throw std::runtime_error("path_simplification \"ompl\" needs OMPL's PathSimplifier::smoothBSpline, which is "
                                     "outside this library; use \"custom\" or \"none\"");
        else {
            std::cerr << "Unknown pruning method" << std::endl;
            throw std::runtime_error("Unknown pruning method");
        }
        for (const auto &wp : pruned) {
            if (!flat.empty() && norm(flat.back(), wp) < 0.05) continue; // :221-224
            flat.push_back(wp);
        }
    }
    return flat;
}

std::vector<epp::Vector3d> PathPlanner::pruneWaypoints(const std::vector<epp::Vector3d> &waypoints) const
{
    if (waypoints.size() < 3) return waypoints;
    std::vector<epp::Vector3d> pruned;
    pruned.push_back(waypoints[0]);
    std::size_t referenceWaypointIdx = 0;
    std::size_t waypointIdx = 2;
    while (waypointIdx < waypoints.size()) { // :241-255, ray checks with canPassGate = true
        if (!worldPtr->checkRayValid(waypoints[referenceWaypointIdx], waypoints[waypointIdx], true)) {
            pruned.push_back(waypoints[waypointIdx - 1]);
            referenceWaypointIdx = waypointIdx - 1;
        }
        waypointIdx++;
    }
    std::cout << "Pruned " << waypoints.size() - pruned.size() << " waypoints from " << waypoints.size() << " to "
              << pruned.size() << std::endl;
    pruned.push_back(waypoints.back());
    return pruned;
}
