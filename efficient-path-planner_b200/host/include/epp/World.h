// World.h — the reference's World (include/World.h:16-127), same class name, constructor and public
// methods, with the OBB registry and R-tree replaced by a GPU-resident world behind the C ABI
// (include/epp_c_api.h).  Single-unit predicates keep their signatures; batched entry points are added.
//
// Behavioural notes:
//  * every predicate runs on the GPU (libepp_b200.so); there is no CPU fallback — without a device the
//    calls throw std::runtime_error carrying epp_last_error();
//  * mutators (resetWorld/addGate/updateGatePosition/addObstacle) edit the host copy; the next check
//    publishes a new immutable device epoch (the reference mutates its R-tree in place, src/World.cpp:57-78);
//  * exceptions carry the reference's messages ("Rotation around x axis is not supported",
//    "Center z position must be zero", ... src/Object.cpp:16-47).
#pragma once

#include <cstdint>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "ConfigParser.h"
#include "LinAlg.h"

struct epp_world; // include/epp_c_api.h

class World {
public:
    explicit World(std::shared_ptr<ConfigParser> configParser);
    ~World();
    World(const World &) = delete;
    World &operator=(const World &) = delete;

    // ---- reference API (include/World.h:29-92) ----------------------------------------------
    void resetWorld();
    void addGate(const int gateId, const epp::VectorXd &coordinates);            // [x,y,z,roll,pitch,yaw,type]
    void updateGatePosition(const int gateId, const epp::VectorXd &coordinates);
    void addObstacle(const int obstacleId, const epp::VectorXd &coordinates);    // [x,y,z,roll,pitch,yaw]
    bool checkPointValidity(const epp::Vector3d &point, const bool canPassGate);
    bool checkPointValidity(const epp::Vector3d &point, const double minDistance) const;
    bool checkRayValid(const epp::Vector3d &start, const epp::Vector3d &end, const bool canPassGate = false);
    void printObbs() const;

    // ---- batched entry points (added) ----------------------------------------------------------
    // Mask convention of the C ABI: bit (i & 31) of word (i >> 5) is 1 iff unit i is valid.
    static std::size_t maskWords(std::size_t n) { return (n + 31) / 32; }
    // xyz: n x 3 row-major
    void checkPointValidityBatch(const double *xyz, std::size_t n, bool canPassGate, uint32_t *maskWords) const;
    void checkPointValidityBatch(const double *xyz, std::size_t n, double minDistance, uint32_t *maskWords) const;
    void checkRayValidBatch(const double *starts, const double *ends, std::size_t n, bool canPassGate,
                            uint32_t *maskWords) const;
    // PathPlanner::checkTrajectoryValidity over n_traj column-major matrices (rows x cols, leading
    // dimension ld, consecutive matrices trajStride doubles apart).
    void checkTrajectoriesBatch(const double *colMajor, std::size_t nTraj, std::size_t rows, std::size_t ld,
                                std::size_t trajStride, int cols, double minDistance, uint32_t *trajValidWords,
                                uint32_t *sampleMaskWords, int64_t *firstInvalidRow) const;

    int size() const;
    uint64_t epoch() const;
    epp_world *handle() const; // commits pending mutations first

private:
    void commitIfDirty() const;
    std::shared_ptr<ConfigParser> configParser;
    epp_world *w = nullptr;
    mutable std::mutex mu;
    mutable bool dirty = true;
};
