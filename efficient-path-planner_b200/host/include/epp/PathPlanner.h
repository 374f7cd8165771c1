// PathPlanner.h — the parts of the reference's PathPlanner that sit on the validity-check path
// (include/PathPlanner.h:15-115): world construction from the nominal gate / obstacle tables, the
// OMPL SpaceInformation with both validators installed, checkTrajectoryValidity (the trajectory-sample
// sweep), updateGatePos and the custom waypoint pruning (ray checks).  planPath is provided by a
// batch-native sampling planner (BatchPlanner.h) instead of the per-call OMPL loop; OMPL itself is
// not re-implemented.
#pragma once

#include <memory>
#include <vector>

#include <ompl/base/SpaceInformation.h>
#include <ompl/base/spaces/RealVectorStateSpace.h>

#include "ConfigParser.h"
#include "LinAlg.h"
#include "MotionValidator.h"
#include "StateValidator.h"
#include "World.h"

class PathPlanner {
public:
    PathPlanner(const epp::MatrixXd &nominalGatePositionAndType, const epp::MatrixXd &nominalObstaclePosition,
                std::shared_ptr<ConfigParser> configParser);

    // src/PathPlanner.cpp:60-78: reset, gates (z forced to 0), obstacles.
    void parseGatesAndObstacles(const epp::MatrixXd &nominalGatePositionAndType,
                                const epp::MatrixXd &nominalObstaclePosition);
    // src/PathPlanner.cpp:80-158 (RRT*/FMT through OMPL in the reference): here a batch planner that
    // validates sample and edge batches through isValidBatch / checkMotionBatch.
    bool planPath(const epp::Vector3d &start, const epp::Vector3d &goal, const double timeLimit,
                  std::vector<epp::Vector3d> &resultPath) const;
    // src/PathPlanner.cpp:170-173
    void updateGatePos(const int gateId, const epp::VectorXd &newPose);
    // src/PathPlanner.cpp:267-280: rows of an N x 10 / N x 11 matrix, position = columns 0, 3, 6.
    bool checkTrajectoryValidity(const epp::MatrixXd &trajectory, const double minDistance) const;
    // Batched form: many trajectories at once, one validity bit each.
    std::vector<bool> checkTrajectoryValidityBatch(const std::vector<epp::MatrixXd> &trajectories,
                                                   const double minDistance) const;
    // src/PathPlanner.cpp:175-230 ("none" and "custom" simplification; "ompl" needs OMPL's PathSimplifier).
    std::vector<epp::Vector3d> includeGates2(std::vector<std::vector<epp::Vector3d>> waypoints) const;

    std::shared_ptr<World> worldPtr;
    const ompl::base::SpaceInformationPtr &spaceInformation() const { return si; }

    // src/PathPlanner.cpp:232-265 (private in the reference; public here so it can be tested directly)
    std::vector<epp::Vector3d> pruneWaypoints(const std::vector<epp::Vector3d> &waypoints) const;

private:
    ompl::base::StateSpacePtr space;
    ompl::base::SpaceInformationPtr si;
    std::shared_ptr<ConfigParser> configParser;
};
