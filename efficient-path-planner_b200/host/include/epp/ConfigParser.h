// ConfigParser.h — same class name and getters as the reference's include/ConfigParserYAML.h:19-68,
// reading the same file through the built-in JSON reader instead of yaml-cpp (absent here).
//
// Differences from the reference parser, both deliberate and both only widening what loads:
//  * an OBB without a "name" key takes its map key as name (the shipped config.json has no names and
//    therefore cannot be loaded by the reference parser, src/ConfigParserYAML.cpp:65);
//  * path_planner_properties / trajectory_generator_properties keys missing from the file take the
//    defaults listed in DESIGN.md (the shipped file lacks 8 + 4 of them).
#pragma once

#include <map>
#include <string>
#include <vector>

#include "JsonMin.h"
#include "Types.h"

class ConfigParser {
public:
    explicit ConfigParser(const std::string &configPath);
    // Parse from JSON text instead of a file (used by the Python binding and tests).
    static ConfigParser fromString(const std::string &jsonText);

    const std::vector<OBBDescription> &getGateGeometryByTypeId(const int typeId) const;
    const std::vector<OBBDescription> &getObstacleGeometry() const;
    const ObjectProperties &getObjectPropertiesByTypeId(const int typeId) const;
    const WorldProperties &getWorldProperties() const;
    const PathPlannerProperties &getPathPlannerProperties() const;
    const TrajectoryGeneratorProperties &getTrajectoryGeneratorProperties() const;

    // gate type ids present in gate_id_to_name_mapping
    std::vector<int> gateTypeIds() const;

private:
    ConfigParser() = default;
    void load(const std::string &text);
    epp::Json config;
    std::map<std::string, std::vector<OBBDescription>> objects;
    std::map<std::string, ObjectProperties> objectProperties;
    WorldProperties worldProperties;
    PathPlannerProperties pathPlannerProperties;
    TrajectoryGeneratorProperties trajectoryGeneratorProperties;

    void parseGeometries(const epp::Json &config);
    void parseObjectProperties(const epp::Json &config);
    void parseWorldProperties(const epp::Json &config);
    void parsePathPlannerProperties(const epp::Json &config);
    void parseTrajectoryGeneratorProperties(const epp::Json &config);
};
