// OnlineTrajGenerator.h — the reference's orchestration class (include/OnlineTrajGenerator.h:19-135,
// src/OnlineTrajGenerator.cpp) above the batched checks: checkpoints either side of every gate, the offline
// pre-computation, the online gate update with its look-ahead validity sweep and re-planning, nearest-time sampling.
// Same constructor and method signatures; every world question goes through World / PathPlanner (GPU), the
// trajectory comes from poly_traj::generateTrajectory (GPU).  Trajectory type "snap" only: "spline" and "optimal"
// throw the reference's "Trajectory type not supported".  The reference re-plans on a detached thread
// (:205-222); here the re-plan runs inside updateGatePos, which is what its Python callers observe after the
// trajectoryCurrentlyUpdating flag clears.
#pragma once

#include <memory>
#include <set>
#include <string>
#include <vector>

#include "ConfigParser.h"
#include "LinAlg.h"
#include "PathPlanner.h"
#include "PathWriter.h"

class OnlineTrajGenerator {
public:
    OnlineTrajGenerator(const epp::Vector3d start, const epp::Vector3d goal, const epp::MatrixXd &nominalGatePositionAndType,
                        const epp::MatrixXd &nominalObstaclePosition, const std::string &configPath);

    void preComputeTraj(const double takeoffTime);
    bool updateGatePos(const int gateId, const epp::VectorXd &newPose, const epp::Vector3d &dronePos,
                       const bool nextGateWithinRange, const double flightTime);
    epp::VectorXd sampleTraj(const double currentTime) const;
    double getTrajEndTime() const;
    epp::MatrixXd getPlannedTraj() const;

    // helpers the tests reach for (private in the reference)
    bool checkGatePassed(const epp::Vector3d &pos1, const epp::Vector3d &pos2, const int gateId) const;
    const std::vector<epp::Vector3d> &getCheckpoints() const { return checkpoints; }
    const std::vector<std::vector<epp::Vector3d>> &getPathSegments() const { return pathSegments; }
    PathPlanner &planner() { return pathPlanner; }

private:
    void getGateCenterAndNormal(const epp::VectorXd &gate, epp::Vector3d &center, epp::Vector3d &normal) const;
    void recomputeTraj(const int gateId, const epp::Vector3d &dronePos, const double flightTime);
    epp::MatrixXd generate(const std::vector<epp::Vector3d> &waypoints, double offset, const epp::Vector3d &vel,
                           const epp::Vector3d &acc) const;

    std::shared_ptr<ConfigParser> configParser;
    epp::MatrixXd nominalGatePositionAndType;
    epp::MatrixXd nominalObstaclePosition;
    PathPlanner pathPlanner;
    PathWriter pathWriter;
    std::vector<epp::Vector3d> checkpoints;
    std::vector<std::vector<epp::Vector3d>> pathSegments;
    epp::MatrixXd plannedTraj;
    std::set<int> gatesObservedWithinRange;
    bool trajectoryCurrentlyUpdating = false;
};
