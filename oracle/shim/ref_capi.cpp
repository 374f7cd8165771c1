// oracle/shim/ref_capi.cpp — TEST INFRASTRUCTURE.  C entry points over the reference's OWN World /
// Object / OBB / StateValidator / MotionValidator, compiled unmodified from /root/reference/src into
// oracle/_ref/libepp_ref.so together with this file (oracle/Makefile target `ref`).  The reference's
// ConfigParser implementation (src/ConfigParserYAML.cpp, needs yaml-cpp) is replaced by the definitions
// below, which fill the same private tables from arrays handed over by the test.
#include <cstdint>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "ConfigParserYAML.h"
#include "MotionValidator.h"
#include "Object.h"
#include "StateValidator.h"
#include "World.h"
#include "PathWriter.h"
#include <ompl/base/spaces/RealVectorStateSpace.h>

namespace ob = ompl::base;

namespace {
struct Spec {
    std::map<int, std::vector<OBBDescription>> gates;
    std::vector<OBBDescription> obstacle;
    double inflateGate = 0.0, inflateObstacle = 0.0;
};
std::map<std::string, Spec> g_specs;
std::mutex g_mu;
} // namespace

// ---- ConfigParser: same class, tables filled from the registry instead of a YAML file ----------------
ConfigParser::ConfigParser(const std::string &configPath)
{
    std::lock_guard<std::mutex> lk(g_mu);
    const Spec &s = g_specs.at(configPath);
    for (const auto &kv : s.gates) objects.insert({"gate_" + std::to_string(kv.first), kv.second});
    objects.insert({"obstacle", s.obstacle});
    for (const auto &kv : s.gates) objectProperties.insert({"gate_" + std::to_string(kv.first), {0.0}});
    worldProperties.inflateRadius["gate"] = s.inflateGate;
    worldProperties.inflateRadius["obstacle"] = s.inflateObstacle;
}
const std::vector<OBBDescription> &ConfigParser::getGateGeometryByTypeId(const int typeId) const
{
    return objects.at("gate_" + std::to_string(typeId));
}
const ObjectProperties &ConfigParser::getObjectPropertiesByTypeId(const int typeId) const
{
    return objectProperties.at("gate_" + std::to_string(typeId));
}
const std::vector<OBBDescription> &ConfigParser::getObstacleGeometry() const { return objects.at("obstacle"); }
const WorldProperties &ConfigParser::getWorldProperties() const { return worldProperties; }
const PathPlannerProperties &ConfigParser::getPathPlannerProperties() const { return pathPlannerProperties; }
const TrajectoryGeneratorProperties &ConfigParser::getTrajectoryGeneratorProperties() const
{
    return trajectoryGeneratorProperties;
}

// ---- C API ------------------------------------------------------------------------------------------------
struct ref_desc { // same layout as orc_obb_desc
    double position[3];
    double size[3];
    int filling;
};

struct ref_world {
    std::string key;
    std::shared_ptr<ConfigParser> cfg;
    std::shared_ptr<World> world;
    ob::SpaceInformationPtr si;
    std::shared_ptr<StateValidator> sv[2];  // canPassGate false / true
    std::shared_ptr<MotionValidator> mv[2];
    Spec spec;
    std::string error;
};

namespace {
std::vector<OBBDescription> descs(const ref_desc *d, int n)
{
    std::vector<OBBDescription> out;
    for (int i = 0; i < n; ++i) {
        // src/ConfigParserYAML.cpp:62-66: halfSize = size / 2
        const Eigen::Vector3d position(d[i].position[0], d[i].position[1], d[i].position[2]);
        const Eigen::Vector3d halfSize = Eigen::Vector3d(d[i].size[0], d[i].size[1], d[i].size[2]) / 2;
        out.push_back({position, halfSize, d[i].filling ? "filling" : "collision", "obb" + std::to_string(i)});
    }
    return out;
}

void build(ref_world *w)
{
    {
        std::lock_guard<std::mutex> lk(g_mu);
        g_specs[w->key] = w->spec;
    }
    w->cfg = std::make_shared<ConfigParser>(w->key);
    w->world = std::make_shared<World>(w->cfg);
    ob::StateSpacePtr space(new ob::RealVectorStateSpace(3));
    w->si = ob::SpaceInformationPtr(new ob::SpaceInformation(space));
    for (int c = 0; c < 2; ++c) {
        w->sv[c] = std::make_shared<StateValidator>(w->si, w->world, c == 1);
        w->mv[c] = std::make_shared<MotionValidator>(w->si, w->world, c == 1);
    }
}

template <class F>
int guarded(ref_world *w, F f)
{
    try {
        f();
        return 0;
    } catch (const std::exception &e) {
        w->error = e.what();
        return 1;
    }
}

Eigen::VectorXd vec(const double *p, int n)
{
    Eigen::VectorXd v(n);
    for (int i = 0; i < n; ++i) v(i) = p[i];
    return v;
}

template <class F>
void parallel(size_t n, int threads, F f)
{
    // whole mask words per thread
    const size_t words = (n + 31) / 32;
    if (threads <= 1 || words < 2) {
        f(0, n);
        return;
    }
    std::vector<std::thread> ts;
    const size_t per = (words + threads - 1) / threads * 32;
    for (size_t b = 0; b < n; b += per) ts.emplace_back(f, b, std::min(n, b + per));
    for (auto &t : ts) t.join();
}
} // namespace

extern "C" {

ref_world *ref_world_create(double inflate_gate, double inflate_obstacle)
{
    static int counter = 0;
    ref_world *w = new ref_world;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        w->key = "mem:" + std::to_string(counter++);
    }
    w->spec.inflateGate = inflate_gate;
    w->spec.inflateObstacle = inflate_obstacle;
    build(w);
    return w;
}
void ref_world_destroy(ref_world *w) { delete w; }
const char *ref_last_error(ref_world *w) { return w->error.c_str(); }

// geometry must be set before objects are added (rebuilds the reference World around a new ConfigParser)
int ref_world_set_gate_geometry(ref_world *w, int type_id, const ref_desc *d, int n)
{
    w->spec.gates[type_id] = descs(d, n);
    build(w);
    return 0;
}
int ref_world_set_obstacle_geometry(ref_world *w, const ref_desc *d, int n)
{
    w->spec.obstacle = descs(d, n);
    build(w);
    return 0;
}
void ref_world_reset(ref_world *w) { w->world->resetWorld(); }
int ref_world_add_gate(ref_world *w, int id, const double *pose7)
{
    return guarded(w, [&] { w->world->addGate(id, vec(pose7, 7)); });
}
int ref_world_update_gate(ref_world *w, int id, const double *pose7)
{
    return guarded(w, [&] { w->world->updateGatePosition(id, vec(pose7, 7)); });
}
int ref_world_add_obstacle(ref_world *w, int id, const double *pose6)
{
    return guarded(w, [&] { w->world->addObstacle(id, vec(pose6, 6)); });
}

// Object::createFromDescription + getAABBs on one object: out[i*24 ..] = center3 half3 rot9(row-major) aabbmin3 aabbmax3
int ref_object_records(ref_world *w, int is_gate, int type_id, const double *pos3, const double *rot3, double inflate,
                       double *out, int max_obbs)
{
    int n = 0;
    const int rc = guarded(w, [&] {
        const std::vector<OBBDescription> &d =
            is_gate ? w->cfg->getGateGeometryByTypeId(type_id) : w->cfg->getObstacleGeometry();
        Object obj = Object::createFromDescription(Eigen::Vector3d(pos3[0], pos3[1], pos3[2]),
                                                   Eigen::Vector3d(rot3[0], rot3[1], rot3[2]), d);
        const std::vector<box> aabbs = obj.getAABBs(inflate);
        for (size_t i = 0; i < obj.obbs.size() && (int)i < max_obbs; ++i, ++n) {
            double *o = out + 24 * i;
            for (int k = 0; k < 3; ++k) {
                o[k] = obj.obbs[i].center(k);
                o[3 + k] = obj.obbs[i].halfSize(k);
                for (int c = 0; c < 3; ++c) o[6 + 3 * k + c] = obj.obbs[i].rotation(k, c);
                o[15 + k] = aabbs[i].min_corner()(k);
                o[18 + k] = aabbs[i].max_corner()(k);
            }
        }
    });
    return rc ? -1 : n;
}

// Batched loops over the reference predicates; `via_validators` routes through StateValidator::isValid /
// MotionValidator::checkMotion on OMPL-shaped states instead of calling World directly.
void ref_check_states(ref_world *w, const double *xyz, size_t n, int can_pass_gate, uint32_t *mask, int threads,
                      int via_validators)
{
    std::memset(mask, 0, ((n + 31) / 32) * 4);
    parallel(n, threads, [&](size_t b, size_t e) {
        ob::RealVectorStateSpace::StateType st;
        double vals[3];
        st.values = vals;
        for (size_t i = b; i < e; ++i) {
            bool ok;
            if (via_validators) {
                std::memcpy(vals, xyz + 3 * i, sizeof vals);
                ok = w->sv[can_pass_gate ? 1 : 0]->isValid(&st);
            } else {
                ok = w->world->checkPointValidity(Eigen::Vector3d(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]),
                                                  can_pass_gate != 0);
            }
            if (ok) mask[i >> 5] |= 1u << (i & 31);
        }
    });
}

void ref_check_states_mindist(ref_world *w, const double *xyz, size_t n, double min_distance, uint32_t *mask, int threads)
{
    std::memset(mask, 0, ((n + 31) / 32) * 4);
    const World &cw = *w->world;
    parallel(n, threads, [&](size_t b, size_t e) {
        for (size_t i = b; i < e; ++i)
            if (cw.checkPointValidity(Eigen::Vector3d(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]), min_distance))
                mask[i >> 5] |= 1u << (i & 31);
    });
}

void ref_check_segments(ref_world *w, const double *a, const double *b, size_t n, int can_pass_gate, uint32_t *mask,
                        int threads, int via_validators)
{
    std::memset(mask, 0, ((n + 31) / 32) * 4);
    parallel(n, threads, [&](size_t lo, size_t hi) {
        ob::RealVectorStateSpace::StateType s1, s2;
        double v1[3], v2[3];
        s1.values = v1;
        s2.values = v2;
        for (size_t i = lo; i < hi; ++i) {
            bool ok;
            if (via_validators) {
                std::memcpy(v1, a + 3 * i, sizeof v1);
                std::memcpy(v2, b + 3 * i, sizeof v2);
                ok = w->mv[can_pass_gate ? 1 : 0]->checkMotion(&s1, &s2);
            } else {
                ok = w->world->checkRayValid(Eigen::Vector3d(a[3 * i], a[3 * i + 1], a[3 * i + 2]),
                                             Eigen::Vector3d(b[3 * i], b[3 * i + 1], b[3 * i + 2]), can_pass_gate != 0);
            }
            if (ok) mask[i >> 5] |= 1u << (i & 31);
        }
    });
}

// ---- PathWriter (src/PathWriter.cpp, compiled unmodified): the debug dumps, for the format tests of N4 ----------
struct ref_writer {
    PathWriter pw;
    explicit ref_writer(const char *folder) : pw(folder) {}
};
ref_writer *ref_writer_create(const char *folder) { return new ref_writer(folder); }
void ref_writer_destroy(ref_writer *w) { delete w; }
static std::vector<Eigen::Vector3d> points_of(const double *xyz, size_t n)
{
    std::vector<Eigen::Vector3d> v;
    for (size_t i = 0; i < n; ++i) v.push_back(Eigen::Vector3d(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]));
    return v;
}
static Eigen::VectorXd vector_of(const double *x, int n)
{
    Eigen::VectorXd v(n);
    for (int i = 0; i < n; ++i) v(i) = x[i];
    return v;
}
void ref_writer_write_path(ref_writer *w, const double *xyz, size_t n) { w->pw.writePath(points_of(xyz, n)); }
void ref_writer_write_checkpoints(ref_writer *w, const double *xyz, size_t n) { w->pw.writeCheckpoints(points_of(xyz, n)); }
void ref_writer_update_gate(ref_writer *w, int id, const double *row, int n) { w->pw.updateGatePos(id, vector_of(row, n)); }
void ref_writer_update_obstacle(ref_writer *w, int id, const double *row, int n) { w->pw.updateObstaclePos(id, vector_of(row, n)); }

} // extern "C"
