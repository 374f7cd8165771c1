// Types.h — configuration structs of the reference (include/Types.h:22-87), same field names.
#pragma once

#include <map>
#include <string>

#include "LinAlg.h"

struct OBBDescription {
    epp::Vector3d center;
    epp::Vector3d halfSize;
    std::string type;
    std::string name;
};

struct ObjectProperties {
    double height;
};

struct WorldProperties {
    epp::Vector3d lowerBound;
    epp::Vector3d upperBound;
    std::map<std::string, double> inflateRadius;
};

struct PathPlannerProperties {
    double optimalityThresholdPercentage;
    double timeLimitOnline;
    double timeLimitOffline;
    double checkpointGateOffset;
    double range;
    double minDistCheckTrajCollision;
    std::string pathSimplification;
    bool recalculateOnline;
    bool advanceForCalculation;
    bool canPassGate;
    std::string planner;
    int samplesFMT;
};

struct TrajectoryGeneratorProperties {
    double maxVelocity;
    double maxAcceleration;
    double samplingInterval;
    std::string type;
    double maxTime;
    double maxTrajDivergence;
    double prependTrajTime;
};
