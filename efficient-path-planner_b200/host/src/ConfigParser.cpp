// ConfigParser.cpp — see ConfigParser.h.  Follows src/ConfigParserYAML.cpp:10-119 of the reference.
#include "epp/ConfigParser.h"

#include <fstream>
#include <sstream>
#include <stdexcept>

using epp::Json;

namespace {

epp::Vector3d vec3(const Json &a) { return epp::Vector3d(a[0].asDouble(), a[1].asDouble(), a[2].asDouble()); }

double numOr(const Json &o, const char *key, double fallback) { return o.has(key) ? o[key].asDouble() : fallback; }
bool boolOr(const Json &o, const char *key, bool fallback) { return o.has(key) ? o[key].asBool() : fallback; }
std::string strOr(const Json &o, const char *key, const char *fallback)
{
    return o.has(key) ? o[key].asString() : std::string(fallback);
}

} // namespace

ConfigParser::ConfigParser(const std::string &configPath)
{
    std::ifstream f(configPath);
    if (!f) throw std::runtime_error("bad file: " + configPath); // YAML::BadFile
    std::stringstream ss;
    ss << f.rdbuf();
    load(ss.str());
}

ConfigParser ConfigParser::fromString(const std::string &jsonText)
{
    ConfigParser p;
    p.load(jsonText);
    return p;
}

void ConfigParser::load(const std::string &text)
{
    config = Json::parse(text);
    parseGeometries(config);
    parseObjectProperties(config);
    parseWorldProperties(config);
    parsePathPlannerProperties(config);
    parseTrajectoryGeneratorProperties(config);
}

const std::vector<OBBDescription> &ConfigParser::getGateGeometryByTypeId(const int typeId) const
{
    const std::string key = config["gate_id_to_name_mapping"][std::to_string(typeId)].asString();
    return objects.at(key);
}

const ObjectProperties &ConfigParser::getObjectPropertiesByTypeId(const int typeId) const
{
    const std::string key = config["gate_id_to_name_mapping"][std::to_string(typeId)].asString();
    return objectProperties.at(key);
}

const std::vector<OBBDescription> &ConfigParser::getObstacleGeometry() const { return objects.at("obstacle"); }
const WorldProperties &ConfigParser::getWorldProperties() const { return worldProperties; }
const PathPlannerProperties &ConfigParser::getPathPlannerProperties() const { return pathPlannerProperties; }
const TrajectoryGeneratorProperties &ConfigParser::getTrajectoryGeneratorProperties() const
{
    return trajectoryGeneratorProperties;
}

std::vector<int> ConfigParser::gateTypeIds() const
{
    std::vector<int> ids;
    for (const auto &kv : config["gate_id_to_name_mapping"].obj) ids.push_back(std::stoi(kv.first));
    return ids;
}

void ConfigParser::parseGeometries(const Json &cfg)
{
    for (const auto &component : cfg["component_geometry"].obj) {
        std::vector<OBBDescription> obbDescriptions;
        for (const auto &obbi : component.second.obj) {
            const Json &obb = obbi.second;
            const epp::Vector3d position = vec3(obb["position"]);
            const epp::Vector3d size = vec3(obb["size"]);
            const epp::Vector3d halfSize(size(0) / 2, size(1) / 2, size(2) / 2); // :63
            const std::string type = obb["type"].asString();
            const std::string name = obb.has("name") ? obb["name"].asString() : obbi.first;
            obbDescriptions.push_back({position, halfSize, type, name});
        }
        objects.insert({component.first, obbDescriptions});
    }
}

void ConfigParser::parseObjectProperties(const Json &cfg)
{
    for (const auto &object : cfg["component_properties"].obj)
        objectProperties.insert({object.first, {object.second["height"].asDouble()}});
}

void ConfigParser::parseWorldProperties(const Json &cfg)
{
    const Json &w = cfg["world_properties"];
    worldProperties.lowerBound = vec3(w["lower_bound"]);
    worldProperties.upperBound = vec3(w["upper_bound"]);
    worldProperties.inflateRadius = std::map<std::string, double>();
    worldProperties.inflateRadius["gate"] = w["inflate_radius"]["gate"].asDouble();
    worldProperties.inflateRadius["obstacle"] = w["inflate_radius"]["obstacle"].asDouble();
}

void ConfigParser::parsePathPlannerProperties(const Json &cfg)
{
    const Json &p = cfg["path_planner_properties"];
    PathPlannerProperties &o = pathPlannerProperties;
    o.optimalityThresholdPercentage = numOr(p, "optimality_threshold_percentage", 0.05);
    o.timeLimitOnline = numOr(p, "time_limit_online", 0.5);
    o.timeLimitOffline = numOr(p, "time_limit_offline", 2.0);
    o.checkpointGateOffset = numOr(p, "checkpoint_gate_offset", 0.21);
    o.range = numOr(p, "range", 0.5);
    o.minDistCheckTrajCollision = numOr(p, "min_dist_check_traj_collision", 0.1);
    o.pathSimplification = strOr(p, "path_simplification", "custom");
    o.recalculateOnline = boolOr(p, "recalculate_online", false);
    o.canPassGate = boolOr(p, "can_pass_gate", false);
    o.advanceForCalculation = boolOr(p, "advance_for_calculation", true);
    o.planner = strOr(p, "planner", "rrt");
    o.samplesFMT = static_cast<int>(numOr(p, "samples_fmt", 1000));
}

void ConfigParser::parseTrajectoryGeneratorProperties(const Json &cfg)
{
    const Json &t = cfg["trajectory_generator_properties"];
    TrajectoryGeneratorProperties &o = trajectoryGeneratorProperties;
    o.maxVelocity = numOr(t, "max_velocity", 1.0);
    o.maxAcceleration = numOr(t, "max_acceleration", 2.0);
    o.samplingInterval = numOr(t, "sampling_interval", 0.1);
    o.type = strOr(t, "type", "snap");
    o.maxTime = numOr(t, "max_time", 15.0);
    o.prependTrajTime = numOr(t, "prepend_traj_time", 1.0);
    o.maxTrajDivergence = numOr(t, "max_traj_divergence", 0.05);
}
