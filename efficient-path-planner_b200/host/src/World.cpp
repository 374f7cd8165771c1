// World.cpp — see World.h.  Thin C++ layer over the C ABI; the arithmetic lives in the CUDA kernels.
#include "epp/World.h"

#include <iostream>
#include <stdexcept>

#include "../../../include/epp_c_api.h"

namespace {

void check(int status)
{
    if (status == EPP_OK) return;
    const std::string msg = epp_last_error();
    if (status == EPP_ERR_NO_GEOMETRY || status == EPP_ERR_NO_SUCH_OBB) throw std::out_of_range(msg); // map::at
    throw std::runtime_error(msg);
}

std::vector<epp_obb_desc> toDescs(const std::vector<OBBDescription> &in)
{
    std::vector<epp_obb_desc> out(in.size());
    for (std::size_t i = 0; i < in.size(); ++i) {
        for (int d = 0; d < 3; ++d) {
            out[i].position[d] = in[i].center(d);
            out[i].size[d] = in[i].halfSize(d) * 2; // exact: the C ABI halves it again (x/2*2 == x for finite x)
        }
        out[i].filling = in[i].type == "filling" ? 1 : 0;
    }
    return out;
}

} // namespace

World::World(std::shared_ptr<ConfigParser> configParser) : configParser(configParser)
{
    const double inflateSizeGate = configParser->getWorldProperties().inflateRadius.at("gate");
    const double inflateSizeObstacle = configParser->getWorldProperties().inflateRadius.at("obstacle");
    check(epp_world_create(inflateSizeGate, inflateSizeObstacle, &w));
    for (int typeId : configParser->gateTypeIds()) {
        const std::vector<epp_obb_desc> d = toDescs(configParser->getGateGeometryByTypeId(typeId));
        check(epp_world_set_gate_geometry(w, typeId, d.data(), static_cast<int>(d.size())));
    }
    const std::vector<epp_obb_desc> d = toDescs(configParser->getObstacleGeometry());
    check(epp_world_set_obstacle_geometry(w, d.data(), static_cast<int>(d.size())));
}

World::~World() { epp_world_destroy(w); }

void World::resetWorld()
{
    std::lock_guard<std::mutex> lk(mu);
    check(epp_world_reset(w));
    dirty = true;
}

void World::addGate(const int gateId, const epp::VectorXd &coordinates)
{
    if (coordinates.size() < 7) throw std::runtime_error("gate coordinates need 7 entries [x,y,z,roll,pitch,yaw,type]");
    double pose[7];
    for (int i = 0; i < 7; ++i) pose[i] = coordinates(i);
    std::lock_guard<std::mutex> lk(mu);
    check(epp_world_add_gate(w, gateId, pose));
    dirty = true;
}

void World::updateGatePosition(const int gateId, const epp::VectorXd &coordinates)
{
    if (coordinates.size() < 7) throw std::runtime_error("gate coordinates need 7 entries [x,y,z,roll,pitch,yaw,type]");
    double pose[7];
    for (int i = 0; i < 7; ++i) pose[i] = coordinates(i);
    std::lock_guard<std::mutex> lk(mu);
    check(epp_world_update_gate(w, gateId, pose));
    dirty = true;
}

void World::addObstacle(const int obstacleId, const epp::VectorXd &coordinates)
{
    if (coordinates.size() < 6) throw std::runtime_error("obstacle coordinates need 6 entries [x,y,z,roll,pitch,yaw]");
    double pose[6];
    for (int i = 0; i < 6; ++i) pose[i] = coordinates(i);
    std::lock_guard<std::mutex> lk(mu);
    check(epp_world_add_obstacle(w, obstacleId, pose));
    dirty = true;
}

void World::commitIfDirty() const
{
    std::lock_guard<std::mutex> lk(mu);
    if (dirty) {
        check(epp_world_commit(w));
        dirty = false;
    }
}

epp_world *World::handle() const
{
    commitIfDirty();
    return w;
}

bool World::checkPointValidity(const epp::Vector3d &point, const bool canPassGate)
{
    commitIfDirty();
    const double p[3] = {point(0), point(1), point(2)};
    int valid = 0;
    check(epp_is_valid(w, p, canPassGate ? 1 : 0, &valid));
    return valid != 0;
}

bool World::checkPointValidity(const epp::Vector3d &point, const double minDistance) const
{
    commitIfDirty();
    const double p[3] = {point(0), point(1), point(2)};
    uint32_t word = 0;
    check(epp_check_states_mindist(w, p, 1, minDistance, &word));
    return (word & 1u) != 0;
}

bool World::checkRayValid(const epp::Vector3d &start, const epp::Vector3d &end, const bool canPassGate)
{
    commitIfDirty();
    const double a[3] = {start(0), start(1), start(2)};
    const double b[3] = {end(0), end(1), end(2)};
    int valid = 0;
    check(epp_check_motion(w, a, b, canPassGate ? 1 : 0, &valid));
    return valid != 0;
}

void World::printObbs() const
{
    int n = 0;
    check(epp_world_size(w, &n));
    for (int i = 0; i < n; ++i) {
        epp_obb o;
        check(epp_world_get_obb(w, i, &o));
        std::cout << "Obb: " << (o.is_gate ? "gate_" : "obstacle_") << o.object_id << "_obb_" << o.obb_index << std::endl;
        std::cout << "Obb center: " << o.center[0] << " " << o.center[1] << " " << o.center[2] << std::endl;
    }
}

void World::checkPointValidityBatch(const double *xyz, std::size_t n, bool canPassGate, uint32_t *maskWords) const
{
    commitIfDirty();
    check(epp_check_states(w, xyz, n, canPassGate ? 1 : 0, maskWords));
}

void World::checkPointValidityBatch(const double *xyz, std::size_t n, double minDistance, uint32_t *maskWords) const
{
    commitIfDirty();
    check(epp_check_states_mindist(w, xyz, n, minDistance, maskWords));
}

void World::checkRayValidBatch(const double *starts, const double *ends, std::size_t n, bool canPassGate,
                               uint32_t *maskWords) const
{
    commitIfDirty();
    check(epp_check_segments(w, starts, ends, n, canPassGate ? 1 : 0, maskWords));
}

void World::checkTrajectoriesBatch(const double *colMajor, std::size_t nTraj, std::size_t rows, std::size_t ld,
                                   std::size_t trajStride, int cols, double minDistance, uint32_t *trajValidWords,
                                   uint32_t *sampleMaskWords, int64_t *firstInvalidRow) const
{
    commitIfDirty();
    check(epp_check_trajectories(w, colMajor, nTraj, rows, ld, trajStride, cols, minDistance, trajValidWords,
                                 sampleMaskWords, firstInvalidRow));
}

int World::size() const
{
    int n = 0;
    check(epp_world_size(w, &n));
    return n;
}

uint64_t World::epoch() const
{
    uint64_t e = 0;
    check(epp_world_epoch(w, &e));
    return e;
}
