// host_selftest — exercises the C++ host layer the way the reference's own code uses its classes.
//   host_selftest <config.json> --no-gpu   API shape, config parsing, error behaviour (no device needed)
//   host_selftest <config.json> --gpu      the KAT answers of SURVEY.md §8c through the OMPL-shaped
//                                          validators, batched == single, trajectory sweep, planner
// Exit code 0 = all checks passed; every failed check prints a line.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <sstream>
#include <stdexcept>

#include "epp/BatchPlanner.h"
#include "epp/PathPlanner.h"
#include "epp/OnlineTrajGenerator.h"
#include "epp/PathWriter.h"
#include "epp/TrajectoryGenerator.h"

namespace ob = ompl::base;

static int failures = 0;
#define CHECK(cond)                                                            \
    do {                                                                       \
        if (!(cond)) {                                                         \
            std::printf("FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond);        \
            ++failures;                                                        \
        }                                                                      \
    } while (0)

#define CHECK_THROWS(expr, text) CHECK(thrown([&] { (void)(expr); }).find(text) != std::string::npos)

template <class F>
static std::string thrown(F f)
{
    try {
        f();
    } catch (const std::exception &e) {
        return e.what();
    }
    return "";
}

static epp::MatrixXd rows(std::initializer_list<std::initializer_list<double>> r)
{
    const std::size_t n = r.size(), c = n ? r.begin()->size() : 0;
    epp::MatrixXd m(n, c);
    std::size_t i = 0;
    for (const auto &row : r) {
        std::size_t j = 0;
        for (double v : row) m(i, j++) = v;
        ++i;
    }
    return m;
}

static void fill(ob::State *s, double x, double y, double z)
{
    double *v = s->as<ob::RealVectorStateSpace::StateType>()->values;
    v[0] = x;
    v[1] = y;
    v[2] = z;
}

static void noGpu(const std::string &cfgPath)
{
    auto cfg = std::make_shared<ConfigParser>(cfgPath);
    CHECK(cfg->getGateGeometryByTypeId(1).size() == 5);
    CHECK(cfg->getGateGeometryByTypeId(0).size() == 5);
    CHECK(cfg->getObstacleGeometry().size() == 1);
    CHECK(cfg->getObstacleGeometry()[0].halfSize(2) == 1.05 / 2);
    CHECK(cfg->getWorldProperties().inflateRadius.at("gate") == 0.2);
    CHECK(cfg->getWorldProperties().lowerBound(0) == -2 && cfg->getWorldProperties().upperBound(2) == 2);
    CHECK(cfg->getObjectPropertiesByTypeId(1).height == 0.525);
    CHECK(cfg->getTrajectoryGeneratorProperties().samplingInterval == 0.1);
    CHECK(!thrown([&] { cfg->getGateGeometryByTypeId(7); }).empty());
    CHECK(!thrown([] { ConfigParser bad("/nonexistent/config.json"); }).empty());

    auto world = std::make_shared<World>(cfg);
    world->addGate(0, epp::VectorXd{0, 0, 0, 0, 0, 0.3, 1});
    world->addObstacle(0, epp::VectorXd{1, 1, 0, 0, 0, 0});
    CHECK(world->size() == 6);
    // src/Object.cpp:16-47 messages
    CHECK(thrown([&] { world->addGate(1, epp::VectorXd{0, 0, 0, 0.1, 0, 0, 1}); }).find("Rotation around x axis is not supported") != std::string::npos);
    CHECK(thrown([&] { world->addGate(1, epp::VectorXd{0, 0, 0, 0, 0.1, 0, 1}); }).find("Rotation around y axis is not supported") != std::string::npos);
    CHECK(thrown([&] { world->addObstacle(1, epp::VectorXd{0, 0, 0.5, 0, 0, 0}); }).find("Center z position must be zero") != std::string::npos);
    CHECK(!thrown([&] { world->addGate(1, epp::VectorXd{0, 0, 0, 0, 0, 0, 9}); }).empty()); // unknown gate type
    CHECK(!thrown([&] { world->updateGatePosition(5, epp::VectorXd{0, 0, 0, 0, 0, 0, 1}); }).empty()); // obbs.at()
    world->resetWorld();
    CHECK(world->size() == 0);

    // validators install into SpaceInformation exactly as src/PathPlanner.cpp:48-51
    ob::StateSpacePtr space(new ob::RealVectorStateSpace(3));
    ob::SpaceInformationPtr si(new ob::SpaceInformation(space));
    si->setStateValidityChecker(std::make_shared<StateValidator>(si, world, false));
    si->setMotionValidator(std::make_shared<MotionValidator>(si, world, false));
    si->setup();
    // three-argument checkMotion: message on stderr, returns true (src/MotionValidator.cpp:19-23)
    ob::State *a = si->allocState(), *b = si->allocState();
    fill(a, 0, 0, 0.4);
    fill(b, 0, 1, 0.4);
    std::pair<ob::State *, double> lastValid(nullptr, 0.0);
    std::stringstream captured;
    std::streambuf *old = std::cerr.rdbuf(captured.rdbuf());
    const bool r3 = si->getMotionValidator()->checkMotion(a, b, lastValid);
    std::cerr.rdbuf(old);
    CHECK(r3 == true);
    CHECK(captured.str() == "Not implemented error\n");
    si->freeState(a);
    si->freeState(b);
}

static void gpu(const std::string &cfgPath)
{
    auto cfg = std::make_shared<ConfigParser>(cfgPath);
    // KAT world: small_portal (type 1) at the origin, yaw 0
    const epp::MatrixXd gates = rows({{0, 0, 0.7 /* forced to 0 */, 0, 0, 0, 1}});
    const epp::MatrixXd obstacles(0, 6);
    PathPlanner planner(gates, obstacles, cfg);
    auto si = planner.spaceInformation();
    struct P { double x, y, z; bool valid; };
    const P pts[] = {{0, 0, 0.525, true}, {0, 0, 0.4, false}, {0.3, 0, 0.525, false}, {0, 0.3, 0.525, true},
                     {1, 1, 1, true},     {0, 0, 0.9, false}, {0, 0, 1.01, true},     {-0.3, 0.1, 0.2, false}};
    std::vector<ob::State *> states;
    for (const P &p : pts) {
        ob::State *s = si->allocState();
        fill(s, p.x, p.y, p.z);
        CHECK(si->isValid(s) == p.valid);
        states.push_back(s);
    }
    const auto *sv = static_cast<const StateValidator *>(si->getStateValidityChecker().get());
    const std::vector<bool> batch = sv->isValidBatch(std::vector<const ob::State *>(states.begin(), states.end()));
    for (std::size_t i = 0; i < states.size(); ++i) CHECK(batch[i] == pts[i].valid);
    struct S { double a[3], b[3]; bool valid; };
    const S segs[] = {{{0, -1, 0.525}, {0, 1, 0.525}, true},   {{0, -1, 0.4}, {0, 1, 0.4}, false},
                      {{1, -1, 1.5}, {1, 1, 1.5}, true},       {{0.6, -1, 0.525}, {0.6, 1, 0.525}, true},
                      {{0, -1, 1.2}, {0, 1, 1.2}, true}};
    const auto *mv = static_cast<const MotionValidator *>(si->getMotionValidator().get());
    std::vector<double> A, B;
    for (const S &s : segs) {
        ob::State *a = si->allocState(), *b = si->allocState();
        fill(a, s.a[0], s.a[1], s.a[2]);
        fill(b, s.b[0], s.b[1], s.b[2]);
        CHECK(si->checkMotion(a, b) == s.valid);
        si->freeState(a);
        si->freeState(b);
        A.insert(A.end(), s.a, s.a + 3);
        B.insert(B.end(), s.b, s.b + 3);
    }
    uint32_t word = 0;
    mv->checkMotionBatch(A.data(), B.data(), 5, &word);
    for (int i = 0; i < 5; ++i) CHECK(((word >> i) & 1u) == (segs[i].valid ? 1u : 0u));
    CHECK((word >> 5) == 0); // tail bits zero
    for (ob::State *s : states) si->freeState(s);

    // World overloads directly (bool vs double resolve as in the reference)
    CHECK(planner.worldPtr->checkPointValidity(epp::Vector3d(0, 0, 0.4), false) == false);
    const World &cw = *planner.worldPtr; // the double overload is the const one (include/World.h:75)
    CHECK(cw.checkPointValidity(epp::Vector3d(0, 0, 0.4), 0.1) == false);
    CHECK(cw.checkPointValidity(epp::Vector3d(0, 0, 0.525), 0.01) == true);

    // trajectory sweep: N x 10, position in columns 0, 3, 6
    epp::MatrixXd traj(50, 10);
    for (int i = 0; i < 50; ++i) {
        traj(i, 0) = 0.0;
        traj(i, 3) = -1.0 + 2.0 * i / 49.0;
        traj(i, 6) = 0.525;
        traj(i, 9) = 0.1 * i;
    }
    CHECK(planner.checkTrajectoryValidity(traj, 0.01) == true); // through the 5 cm slot with a 1 cm margin
    for (int i = 0; i < 50; ++i) traj(i, 6) = 0.275;
    CHECK(planner.checkTrajectoryValidity(traj, 0.01) == false); // through the bottom bar (z 0.25..0.30)
    CHECK(planner.checkTrajectoryValidity(epp::MatrixXd(0, 10), 0.1) == true);
    epp::MatrixXd high = traj;
    for (int i = 0; i < 50; ++i) high(i, 6) = 1.8;
    const std::vector<bool> tb = planner.checkTrajectoryValidityBatch({traj, high, traj}, 0.1);
    CHECK(tb.size() == 3 && !tb[0] && tb[1] && !tb[2]);

    // gate update publishes a new epoch and changes answers
    const uint64_t e0 = planner.worldPtr->epoch();
    planner.updateGatePos(0, epp::VectorXd{1.0, 1.0, 0, 0, 0, 0, 1});
    CHECK(planner.worldPtr->checkPointValidity(epp::Vector3d(0, 0, 0.4), false) == true);
    CHECK(planner.worldPtr->checkPointValidity(epp::Vector3d(1, 1, 0.4), false) == false);
    CHECK(planner.worldPtr->epoch() == e0 + 1);

    // batch planner on the track world: path found, every vertex and edge valid, ends at the goal
    const epp::MatrixXd tg = rows({{0.45, -1.0, 0, 0, 0, 2.35, 1}, {1.0, -1.55, 0, 0, 0, -0.78, 0},
                                   {0.0, 0.5, 0, 0, 0, 0, 1},      {-0.5, -0.5, 0, 0, 0, 3.14, 0}});
    const epp::MatrixXd to = rows({{1.0, -0.5, 0, 0, 0, 0}, {0.5, -1.5, 0, 0, 0, 0}, {-0.5, 0, 0, 0, 0, 0}, {0, 1.0, 0, 0, 0, 0}});
    PathPlanner track(tg, to, cfg);
    std::vector<epp::Vector3d> path;
    const bool solved = track.planPath(epp::Vector3d(1, 1, 0.05), epp::Vector3d(0, -2, 0.5), 2.0, path);
    CHECK(solved);
    if (solved) {
        CHECK(path.size() >= 2);
        CHECK(path.front()(0) == 1 && path.front()(2) == 0.05);
        CHECK(path.back()(1) == -2 && path.back()(2) == 0.5);
        for (std::size_t i = 0; i < path.size(); ++i) CHECK(track.worldPtr->checkPointValidity(path[i], false));
        for (std::size_t i = 1; i < path.size(); ++i) CHECK(track.worldPtr->checkRayValid(path[i - 1], path[i], false));
        std::printf("planned path: %zu vertices\n", path.size());
    }
    const std::vector<epp::Vector3d> pruned = track.pruneWaypoints(path);
    CHECK(pruned.size() <= path.size() && pruned.size() >= 2);

    // poly_traj::generateTrajectory through the C ABI: interpolation, start state, rest at the end, time column
    {
        const std::vector<epp::Vector3d> wp = {epp::Vector3d(0, 0, 1), epp::Vector3d(1, 0.5, 1.2), epp::Vector3d(1.5, -0.5, 0.8),
                                               epp::Vector3d(0.2, -1.0, 1.0)};
        epp::MatrixXd traj;
        CHECK(poly_traj::generateTrajectory(wp, 2.0, 2.0, 0.05, 0.5, epp::Vector3d(0.1, 0, 0), epp::Vector3d(0, 0, 0), traj));
        CHECK(traj.cols() == 10 && traj.rows() > 20);
        CHECK(traj(0, 9) == 0.5 && traj(0, 0) == 0.0 && traj(0, 6) == 1.0 && std::fabs(traj(0, 1) - 0.1) < 1e-12);
        const std::size_t e = traj.rows() - 1;
        CHECK(std::fabs(traj(e, 0) - 0.2) < 0.05 && std::fabs(traj(e, 3) + 1.0) < 0.05 && std::fabs(traj(e, 1)) < 0.2);
        for (std::size_t r = 1; r < (std::size_t)traj.rows(); ++r) CHECK(std::fabs(traj(r, 9) - traj(r - 1, 9) - 0.05) < 1e-9);
        CHECK_THROWS(poly_traj::generateTrajectory({wp[0]}, 2.0, 2.0, 0.05, 0.0, wp[0], wp[0], traj), "At least two waypoints");
    }

    // OnlineTrajGenerator on the track of config.json: pre-compute, sample, gate update with re-plan
    {
        OnlineTrajGenerator gen(epp::Vector3d(1, 1, 0.05), epp::Vector3d(0, -2, 0.5), tg, to, cfgPath);
        CHECK_THROWS(gen.sampleTraj(0.0), "No trajectory data available.");
        CHECK_THROWS(gen.getTrajEndTime(), "No trajectory data available.");
        CHECK(gen.getCheckpoints().size() == 2 * 4 + 2);
        gen.preComputeTraj(0.5);
        const epp::MatrixXd traj = gen.getPlannedTraj();
        CHECK(traj.cols() == 10 && traj.rows() > 20 && traj(0, 9) == 0.5);
        CHECK(gen.getTrajEndTime() == traj(traj.rows() - 1, 9));
        CHECK(gen.getPathSegments().size() == 5);
        for (const std::vector<epp::Vector3d> &seg : gen.getPathSegments())
            for (std::size_t i = 1; i < seg.size(); ++i) CHECK(gen.planner().worldPtr->checkRayValid(seg[i - 1], seg[i], false));
        auto pos = [](const epp::MatrixXd &t, std::size_t r) { return epp::Vector3d(t(r, 0), t(r, 3), t(r, 6)); };
        auto gatesInOrder = [&](const epp::MatrixXd &t) {
            std::size_t row = 0;
            for (int g = 0; g < 4; ++g) {
                while (row + 1 < (std::size_t)t.rows() && !gen.checkGatePassed(pos(t, row), pos(t, row + 1), g)) ++row;
                if (row + 1 >= (std::size_t)t.rows()) return false;
            }
            return true;
        };
        CHECK(gatesInOrder(traj));
        const std::size_t mid = traj.rows() / 2;
        const epp::VectorXd s = gen.sampleTraj(traj(mid, 9) + 0.01);
        CHECK(s.size() == 10 && s(9) == traj(mid, 9) && s(0) == traj(mid, 0));
        // just after gate 0 the drone sees gate 1 fifteen centimetres off its nominal place
        std::size_t r0 = 0;
        while (r0 + 1 < (std::size_t)traj.rows() && !gen.checkGatePassed(pos(traj, r0), pos(traj, r0 + 1), 0)) ++r0;
        r0 += 2;
        epp::VectorXd pose{1.0 + 0.12, -1.55 + 0.09, 0, 0, 0, -0.78};
        CHECK(gen.updateGatePos(1, pose, pos(traj, r0), false, traj(r0, 9)) == false);
        CHECK(gen.updateGatePos(1, pose, pos(traj, r0), true, traj(r0, 9)) == true);
        const epp::MatrixXd moved = gen.getPlannedTraj();
        bool kept = true;
        for (std::size_t r = 0; r <= r0; ++r)
            for (std::size_t c = 0; c < 10; ++c) kept = kept && moved(r, c) == traj(r, c);
        CHECK(kept);
        CHECK(gatesInOrder(moved));
        CHECK(gen.updateGatePos(1, pose, pos(traj, r0), true, traj(r0, 9)) == false); // seen once: ignored
        std::printf("generator: %zu rows, %zu after the gate update\n", (std::size_t)traj.rows(), (std::size_t)moved.rows());
    }
}

// PathWriter dumps of a fixed call sequence; tests/test_host_cpp.py compares the files byte for byte with the Python
// PathWriter's (which is itself checked against the reference's own PathWriter.cpp, tests/test_path_writer.py).
static int writerDumps(const char *folder)
{
    PathWriter w(folder);
    const double gates[2][7] = {{0.45, -1.0, 0, 0, 0, 2.35, 1}, {1.0, -1.55, 0, 0, 0, -0.78, 0}};
    const double obstacle[6] = {1.0, -0.5, 0, 0, 0, 0};
    w.updateGatePos(0, epp::VectorXd(gates[0], 7));
    w.updateGatePos(1, epp::VectorXd(gates[1], 7));
    w.updateObstaclePos(0, epp::VectorXd(obstacle, 6));
    std::vector<epp::Vector3d> pts = {epp::Vector3d(1.0, 1.0, 0.05), epp::Vector3d(1e-7, -123456.789, 1e6),
                                      epp::Vector3d(1.0 / 3.0, 2e10, -0.0), epp::Vector3d(-0.5, 0.29, 1.0000004)};
    w.writeCheckpoints(pts);
    w.writePath(pts);
    pts.pop_back();
    w.writePath(pts);
    const double moved[7] = {0.123456789, -0.09, 0, 0, 0, 1e-3, 1};
    w.updateGatePos(1, epp::VectorXd(moved, 7));
    return 0;
}

int main(int argc, char **argv)
{
    if (argc == 3 && !std::strcmp(argv[1], "--writer")) return writerDumps(argv[2]);
    if (argc < 3) {
        std::fprintf(stderr, "usage: %s <config.json> --no-gpu|--gpu\n", argv[0]);
        return 2;
    }
    try {
        if (!std::strcmp(argv[2], "--gpu"))
            gpu(argv[1]);
        else
            noGpu(argv[1]);
    } catch (const std::exception &e) {
        std::printf("FAIL uncaught exception: %s\n", e.what());
        ++failures;
    }
    std::printf(failures ? "host_selftest: %d failure(s)\n" : "host_selftest: ok\n", failures);
    return failures ? 1 : 0;
}
